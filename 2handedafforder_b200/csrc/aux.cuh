// Launchers of aux.cu: on-device frame preprocessing, [SEG] row projection, mask evaluation metrics.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "kernels.cuh"

namespace bsam {

constexpr int kResamplePrecisionBits = 32 - 8 - 2;   // Pillow PRECISION_BITS for 8-bit channels

// One axis of Pillow's separable resample: window start / length / 22-bit fixed-point weights per output index.
struct ResampleAxis {
  int in_size = 0, out_size = 0, ksize = 0;
  bool identity = true;
  std::vector<int> lo, cnt, kk;
};
void build_resample_axis(int in_size, int out_size, ResampleAxis* ax);
// most input rows any 8-row output tile of preprocess_kernel needs (sizes its shared-memory tile)
int preprocess_tile_rows(const ResampleAxis& ay, int new_h);
// second-generation kernel: widest input pixel span of a 128-column output tile, most input rows of a 16-row tile
void preprocess_lut(float* out768);
void preprocess2_extents(const ResampleAxis& ax, const ResampleAxis& ay, int new_w, int new_h, int* span_px, int* rows);

// Device copy of the two axes for one (H, W) -> (new_h, new_w); cached in the handle per input size.
struct ResampleTables {
  int H = 0, W = 0, new_h = 0, new_w = 0, kx = 0, ky = 0, ident_x = 1, ident_y = 1, tile_rows = 0;
  int span_px128 = 0, rows16 = 0;    // preprocess2_extents
  float* d_lut = nullptr;            // preprocess_lut on the device
  int *d_xlo = nullptr, *d_xcnt = nullptr, *d_xkk = nullptr, *d_ylo = nullptr, *d_ycnt = nullptr, *d_ykk = nullptr;
};

cudaError_t launch_preprocess(const uint8_t* frames, int B, int H, int W, const ResampleTables& t, void* out,
                              int out_dtype, cudaStream_t s);
cudaError_t launch_seg_scan(const int64_t* ids, int b, int T, int L, int64_t seg_idx, int* rows, int max_rows,
                            int* counts, cudaStream_t s);
cudaError_t launch_rows_fc(const void* x, int x_dtype, long x_ld, const int* rows, const int* n_ptr, int max_rows,
                           const void* Wt, const void* bias, int w_dtype, int K, int N, bool relu, float* y,
                           int n_sm, cudaStream_t s);
cudaError_t launch_mask_metrics(const uint8_t* left, const uint8_t* right, const float* taxonomy, const uint8_t* gt,
                                const int* gt_index, int P, int H, int W, int64_t* counts, uint8_t* pred_out,
                                int n_sm, cudaStream_t s);

}  // namespace bsam
