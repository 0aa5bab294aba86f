#!/usr/bin/env python
"""Phase timeline of CTA 0 of the global attention kernel (tracing variant 2): the first 16 key tiles of its second work
item.  Softmax warps 4 (M tile 0) and 8 (M tile 1): s_full seen / scores loaded / max+rescale done / exp issued /
p_full arrived; MMA warp: per tile t: entered, p_full seen, P.V + S issued."""
import ctypes as C
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

lib = importlib.import_module("2handedafforder_b200._lib").load()
S, nb, heads, hd = 64, 8, 16, 80
g = torch.Generator().manual_seed(94)
rows = nb * S * S
qkv = (torch.randn((rows, 3 * heads * hd), generator=g) * 1.5).to(torch.bfloat16).cuda()
rel_h = (torch.randn((2 * S - 1, hd), generator=g) * 0.3).cuda()
rel_w = (torch.randn((2 * S - 1, hd), generator=g) * 0.3).cuda()
out = torch.zeros((rows, heads * hd), dtype=torch.bfloat16, device="cuda")
trace = torch.zeros((20, 16, 10), dtype=torch.int64, device="cuda")
assert lib.b200sam_debug_set_trace(C.c_void_p(trace.data_ptr())) == 0
import sys as _s
V = int(_s.argv[1]) if len(_s.argv) > 1 else 2
assert lib.b200sam_debug_set_variant(1, V) == 0
ms = C.c_float()
rc = lib.b200sam_debug_attention_bench(C.c_void_p(qkv.data_ptr()), nb, S, heads, hd, C.c_void_p(rel_h.data_ptr()),
                                       C.c_void_p(rel_w.data_ptr()), C.c_void_p(out.data_ptr()), 1, C.byref(ms),
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream))
torch.cuda.synchronize()
lib.b200sam_debug_set_variant(1, 0)
lib.b200sam_debug_set_trace(None)
assert rc == 0
t = trace.cpu().numpy()
t0 = int(t[t > 0].min())
rel = lambda v: int(v - t0) if v > 0 else -1
print(f"ms per launch (G GEMM + traced kernel) {ms.value:.4f}")
print("MMA warp per key tile j: tile0: enter, p_full seen, issued | tile1: enter, p_full seen, issued")
for j in range(16):
    print(f"  j{j:2d}: " + " ".join(f"{rel(t[1, j, k]):7d}" for k in (0, 1, 2)) + " | " +
          " ".join(f"{rel(t[1, j, k]):7d}" for k in (4, 5, 6)))
for w in (4, 8):
    print(f"softmax warp {w}: s_full seen, then deltas: load, max/rescale, exp, arrive; wait for next S; period")
    for j in range(1, 15):
        m = [rel(t[w, j, k]) for k in range(5)]
        d = [m[k] - m[k - 1] for k in range(1, 5)]
        print(f"  j{j:2d} @{m[0]:7d}  load={d[0]} max={d[1]} exp={d[2]} arrive={d[3]}  wait_S={rel(t[w, j + 1, 0]) - m[4]}"
              f"  period={rel(t[w, j + 1, 0]) - m[0]}")
