#!/usr/bin/env python
"""bench.py - frames/s of ViT-H encode + bimanual mask decode at 1024^2 on N B200 (BASELINE.json metric).

One "step" = one `visual_model.segment(...)` call over a batch of B synthetic frames x 3 [SEG] prompts:
ImageEncoderViT (ViT-H) -> PromptEncoder text path -> mask_decoder_left + mask_decoder_right -> postprocess_masks to
1080x1920 -> threshold (u8 masks).  N > 1: every rank runs its own frames (weak scaling, frames are independent:
SURVEY.md 8e) and the u8 masks are gathered to rank 0 over NCCL inside the step.

    python bench.py --gpus 1 --steps 10 --warmup 3              # this repo's CUDA path
    python bench.py --impl reference --steps 1 --warmup 0       # CPU arm: the oracle port on the host cores

JSON keys follow the driver contract: value (device-resident inputs, CUDA events, max over ranks), e2e (pinned
host inputs, H2D + D2H inside the timed region), roofline (dominant kernel class, measured with per-launch CUDA
events through b200sam_profile_*), cpu_baseline (oracle port on the host cores, bounded sample), clocks,
gpu_launches.  Nothing here reads /root/reference.
"""
from __future__ import annotations

import argparse
import ctypes as C
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG = "2handedafforder_b200"
IMG, GRID, OUT_CH = 1024, 64, 256
RESIZE, ORIGINAL = (576, 1024), (1080, 1920)      # 16:9 video frame, inference.py:91-105
VIT_H = dict(dim=1280, depth=32, heads=16, global_idx=(7, 15, 23, 31))

PC_NAMES = ["elementwise", "gemm_patch", "layernorm", "gemm_qkv", "attn_window", "attn_global", "gemm_proj",
            "gemm_lin1", "gemm_lin2", "neck", "dec_elementwise", "dec_token", "dec_gemm", "dec_upscale",
            "postprocess", "preprocess", "seg_project", "metrics"]


def flops_per_frame(n_prompts: int) -> dict:
    """Algorithmic FLOPs per frame by kernel class (multiply-add = 2, unpadded tokens)."""
    D, T, depth, hd = VIT_H["dim"], GRID * GRID, VIT_H["depth"], VIT_H["dim"] // VIT_H["heads"]
    n_glob = len(VIT_H["global_idx"])
    n_win = depth - n_glob
    f = {
        "gemm_patch": 2.0 * T * 768 * D,
        "gemm_qkv": depth * 2.0 * T * D * 3 * D,
        "gemm_proj": depth * 2.0 * T * D * D,
        "gemm_lin1": depth * 2.0 * T * D * 4 * D,
        "gemm_lin2": depth * 2.0 * T * D * 4 * D,
        "attn_window": n_win * (2.0 * 2 * T * 196 * D + 2.0 * 2 * T * 14 * D),
        "attn_global": n_glob * (2.0 * 2 * T * T * D + 2.0 * 2 * T * 64 * D),
        "neck": 2.0 * T * D * OUT_CH + 2.0 * T * OUT_CH * OUT_CH * 9,
        "decoder": n_prompts * 2 * 3.606e9 + n_prompts * 0.004e9,
    }
    f["total"] = sum(f.values())
    return f


def load_peaks() -> dict:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            d = json.load(fh)
        return {"bf16_burst": d["bf16_tflops"], "bf16_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "hbm_gbs": d["hbm_gbs"], "source": "measured (MEASURED_PEAKS.json)"}
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def ncu_traffic(kernel_class):
    """DRAM bytes per launch of a kernel class from the committed `ncu --set full` capture (profiles/ncu_traffic.json,
    written by scripts/ncu_summarise.py), or None."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None
    with open(p) as fh:
        return json.load(fh).get(kernel_class)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.t.join(timeout=2)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # under load = samples whose power draw is above half the max seen
        thr = 0.5 * max(pw)
        load = [s for s, p in zip(sm, pw) if p >= thr] or sm
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------- synthetic data

def make_frames(torch, start: int, count: int):
    """(count,3,1024,1024) fp32 ~ N(0,1), rows >= 576 zero (16:9 frame padded to square); seed 10+index."""
    out = torch.empty((count, 3, IMG, IMG), dtype=torch.float32)
    for i in range(count):
        g = torch.Generator().manual_seed(10 + start + i)
        out[i] = torch.randn(3, IMG, IMG, generator=g)
    out[:, :, RESIZE[0]:, :] = 0.0
    return out


def make_seg(torch, index: int, n: int):
    g = torch.Generator().manual_seed(100003 + index)
    return torch.randn(n, OUT_CH, generator=g)


def build_model(torch, max_batch: int, max_prompts: int):
    """Random-init ViT-H visual_model (the facade's own seeded init; pos_embed / rel_pos_* get N(0, 0.02) noise so the
    rel-pos path does real arithmetic - they are zeros at init in the reference, image_encoder.py:70-74,232-233)."""
    pkg = importlib.import_module(PKG)
    sam = pkg.build_sam_vit_h(max_batch=max_batch, max_prompts=max_prompts)
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for name, p in sam.named_parameters():
            if "pos_embed" in name or "rel_pos" in name:
                p.copy_(0.02 * torch.randn(p.shape, generator=g))
    return sam


# ----------------------------------------------------------------------------------------------- CPU arm

def cpu_reference_frames_per_s(torch, sd, n_prompts: int, steps: int, warmup: int, arch_name: str = "vit_h"):
    """The oracle port (oracle/sam_oracle.py, fp32, all host threads) on ONE frame x n_prompts per step:
    encode + both decoders + postprocess + threshold.  Returns (frames/s, seconds per step list, threads)."""
    from oracle import sam_oracle as O
    from oracle import synth
    arch = synth.ARCHS[arch_name]
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    times = []
    with torch.no_grad():
        for it in range(warmup + steps):
            frames = make_frames(torch, it, 1)
            segs = [make_seg(torch, it, n_prompts)]
            t0 = time.perf_counter()
            res = O.segment(sd, arch, frames, segs, [RESIZE], [ORIGINAL], O.FP32)[0]
            for side in ("left", "right"):
                _ = (res[f"masks_{side}"] > 0).to(torch.uint8)
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append(dt)
    return len(times) / sum(times), times, threads


def cpu_baseline_block(torch, sd_vit_h, n_prompts: int) -> dict:
    """the CPU path timed with 1 warm-up + median of 3 on (a) the headline workload's model,
    ViT-H 1 frame x n_prompts, and (b) BASELINE config 1 exactly (ViT-B, 1 frame, 2 prompts).  `value` is (a)."""
    from oracle import synth
    _, t_h, threads = cpu_reference_frames_per_s(torch, sd_vit_h, n_prompts, 3, 1, "vit_h")
    sd_b = synth.make_state_dict(synth.ARCHS["vit_b"], seed=0)
    _, t_b, _ = cpu_reference_frames_per_s(torch, sd_b, 2, 3, 1, "vit_b")
    med_h, med_b = statistics.median(t_h), statistics.median(t_b)
    return {"value": 1.0 / med_h, "unit": "frames/s", "cores": threads, "kind": "port",
            "sample": f"1 ViT-H frame x {n_prompts} prompts, fp32 oracle port, 1 warm-up + median of 3 "
                      f"({med_h:.2f} s per frame)",
            "config1_vit_b": {"value": 1.0 / med_b, "unit": "frames/s", "seconds_per_frame": med_b,
                              "sample": "BASELINE configs[0]: ViT-B, 1 frame x 2 prompts, fp32, 1 warm-up + median of 3"}}


def run_reference_arm(args):
    """--impl reference: the reference's CPU path (oracle port; the reference is Python and cannot travel to the GPU
    box) timed on the host cores.  Rank 0 only; other ranks exit 0."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    import torch
    sam = build_model(torch, 1, args.prompts)
    sd = {k: v.detach().float() for k, v in sam.state_dict().items()}
    steps = max(1, args.steps)
    fps, times, threads = cpu_reference_frames_per_s(torch, sd, args.prompts, steps, args.warmup)
    sample = f"1 ViT-H frame x {args.prompts} prompts per step (encode + 2 decoders + postprocess 1080x1920), fp32"
    line = {
        "impl": "reference", "metric": "frames/s ViT-H encode + bimanual mask decode @1024^2", "value": fps,
        "unit": "frames/s", "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, frames_per_step=1),
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_line(json.dumps(line))
    return 0


def workload_config(args, frames_per_step: int) -> dict:
    return {"workload": "configs[1]: SAM ViT-H bf16, batch of synthetic 1024^2 frames x 3 [SEG] prompts "
                        "(left+right decoders, taxonomy), postprocess to 1080x1920 + threshold, masks in the 1-bit "
                        "wire format (numpy.packbits order)",
            "frames_per_step_per_gpu": frames_per_step, "prompts_per_frame": args.prompts,
            "resize": list(RESIZE), "original_size": list(ORIGINAL), "parallelism": f"frames sharded x{args.gpus}",
            "l2": "per-step working set (1.27 GB bf16 weights + >4 GB activations) >> 126 MB L2; no explicit flush"}


# ----------------------------------------------------------------------------------------------- GPU arm

def measure_callers(torch, sam, dev, last_out, B, NP, iters=20):
    """The HBM-bound kernels either side of the branch, each timed alone with CUDA events
    and set against the measured copy bandwidth with ALGORITHMIC bytes: preprocess = raw u8 frame in + bf16 1024^2
    tensor out; seg_project = one read of the two weight matrices; mask_metrics = three u8 masks in."""
    from torch import nn
    peaks = load_peaks()
    H, W = ORIGINAL

    _lib = importlib.import_module(PKG + "._lib")
    lib, handle = _lib.load(), sam._handle

    def timed(fn, cls):
        """(ms per call through the Python API, ms of this library's kernels alone per call: CUDA events around
        every launch of the class, b200sam_profile_*)."""
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        api_ms = e0.elapsed_time(e1) / iters
        lib.b200sam_profile_enable(handle, 1)
        for _ in range(iters):
            fn()
        torch.cuda.synchronize()
        prof = read_profile(lib, handle)
        lib.b200sam_profile_enable(handle, 0)
        return api_ms, prof[cls][0] / iters

    def entry(api_ms, k_ms, by, what):
        return {"ms": k_ms, "api_ms": api_ms, "bytes": by, "gbs": by / k_ms / 1e6,
                "frac_hbm": by / k_ms / 1e6 / peaks["hbm_gbs"], "what": what}

    res = {}
    g = torch.Generator().manual_seed(77)
    raw = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g).to(dev)
    a, k = timed(lambda: sam.preprocess_frames(raw, torch.bfloat16), "preprocess")
    res["preprocess"] = entry(a, k, B * (H * W * 3 + 3 * IMG * IMG * 2),
                              f"{B} raw {H}x{W} RGB u8 frames -> (B,3,1024,1024) bf16")
    hdim, T, seg = 4096, 64, 32003
    fcs = nn.ModuleList([nn.Sequential(nn.Linear(hdim, hdim), nn.ReLU(), nn.Linear(hdim, OUT_CH), nn.Dropout(0.0))]
                        ).to(torch.bfloat16).to(dev)
    ids = torch.randint(0, 32000, (B, T), generator=g)
    ids[:, T - NP:] = seg
    ids = ids.to(dev)
    hidden = torch.randn((B, 255 + T - 1, hdim), generator=g).to(torch.bfloat16).to(dev)
    a, k = timed(lambda: sam.seg_project(hidden, ids, seg, fcs, max_prompts=B * NP), "seg_project")
    res["seg_project"] = entry(a, k, (hdim * hdim + OUT_CH * hdim) * 2,
                               f"{B * NP} [SEG] rows, LLaVA-7B hidden {hdim}, bf16 weights (scan + 2 FC kernels; "
                               "api_ms includes the count read-back)")
    ml, mr, tx = last_out["masks_left_packed"], last_out["masks_right_packed"], last_out["taxonomy"]
    gt = (ml | mr).contiguous()
    a, k = timed(lambda: sam.mask_metrics(ml, mr, tx, gt), "metrics")
    res["mask_metrics"] = entry(a, k, 3 * ml.numel(), f"{ml.shape[0]} prompts x {H}x{W} u8 masks (left, right, gt)")
    return res


def measure_latency_b1(torch, sam, frame, seg, iters: int = 10):
    """One frame x its prompts through the drop-in exactly as LISAForCausalLM.evaluate drives the reference modules
    (LISA.py:162 image_encoder, :498-503 prompt_encoder, :505-530 both decoders, :511-527 postprocess_masks), then the
    callers' threshold and the copy of the masks to the host (chat.py:231, inference.py:276-301).  CUDA-event time of
    the device work and wall time including the host read-back, median over `iters` after 2 warm-ups."""
    ev_ms, wall_ms = [], []
    with torch.no_grad():
        for it in range(iters + 2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            emb = sam.image_encoder(frame)
            sparse, dense = sam.prompt_encoder(points=None, boxes=None, masks=None, text_embeds=seg.unsqueeze(1))
            sparse = sparse.to(seg.dtype)
            pe = sam.prompt_encoder.get_dense_pe()
            low_l, iou_l, tax = sam.mask_decoder_left(image_embeddings=emb[0].unsqueeze(0), image_pe=pe,
                                                      sparse_prompt_embeddings=sparse,
                                                      dense_prompt_embeddings=dense, multimask_output=False)
            pl = sam.postprocess_masks(low_l, input_size=RESIZE, original_size=ORIGINAL)
            low_r, iou_r = sam.mask_decoder_right(image_embeddings=emb[0].unsqueeze(0), image_pe=pe,
                                                  sparse_prompt_embeddings=sparse, dense_prompt_embeddings=dense,
                                                  multimask_output=False)
            pr = sam.postprocess_masks(low_r, input_size=RESIZE, original_size=ORIGINAL)
            ml, mr = pl[:, 0] > 0, pr[:, 0] > 0
            e1.record()
            _ = ml.cpu(), mr.cpu(), tax.cpu()
            wall = (time.perf_counter() - t0) * 1e3
            if it >= 2:
                ev_ms.append(e0.elapsed_time(e1))
                wall_ms.append(wall)
    return {"device_ms": statistics.median(ev_ms), "wall_ms_incl_host_masks": statistics.median(wall_ms),
            "what": f"1 ViT-H frame x {seg.shape[0]} prompts via image_encoder / prompt_encoder / mask_decoder_left / "
                    "mask_decoder_right / postprocess_masks (fp32 logits) as LISA.evaluate calls them, bool masks to host"}


def read_profile(lib, handle):
    ms = (C.c_float * len(PC_NAMES))()
    cnt = (C.c_int64 * len(PC_NAMES))()
    rc = lib.b200sam_profile_read(handle, ms, cnt, len(PC_NAMES))
    if rc != 0:
        raise RuntimeError(f"b200sam_profile_read status {rc}")
    return {n: (float(ms[i]), int(cnt[i])) for i, n in enumerate(PC_NAMES)}


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (sm_100a): libb200sam has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # keep stdout to the one JSON line: NCCL prints its version banner there at NCCL_DEBUG=VERSION
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world

    B, NP = args.batch, args.prompts
    P = B * NP
    _lib = importlib.import_module(PKG + "._lib")
    lib = _lib.load()
    sam = build_model(torch, B, P).to(dev)
    sam.freeze_weights(True)

    # this rank's shard of the synthetic clip: frames [rank*B, rank*B + B)
    host_frames = make_frames(torch, rank * B, B).to(torch.bfloat16).pin_memory()
    host_seg = torch.cat([make_seg(torch, rank * B + i, NP) for i in range(B)], 0).pin_memory()
    frames = host_frames.to(dev)
    segs = [host_seg[i * NP:(i + 1) * NP].to(dev) for i in range(B)]
    resize, original = [RESIZE] * B, [ORIGINAL] * B
    H, W = ORIGINAL
    clip = importlib.import_module(PKG + ".clip")

    lay = clip.SlotLayout(B, NP, ORIGINAL)
    slot_dev = torch.zeros((lay.nbytes,), dtype=torch.uint8, device=dev)     # N = 1: the message the kernels write
    slot_bits, slot_heads, _, _ = lay.views(slot_dev)

    def step(fr, sg):
        # masks leave the device in the 1-bit wire format (b200sam_postprocess_pair): same information as the u8
        # masks of segment(), an eighth of the bytes
        return sam.segment_bits(fr, sg, RESIZE, ORIGINAL, threshold=0.0, bits_out=slot_bits, heads_out=slot_heads)

    clip_stats = {"masks_checked": 0, "checksum_mismatches": 0, "frames": 0}

    def run_steps(k, load, batch_sink=None, prefetch=False):
        """k steps.  N = 1: k segment_bits() calls.  N > 1: the product clip path - a clip of world*B*k frames sharded
        over the ranks, one micro-batch of B frames per rank per step, packed masks + heads + checksums gathered to
        rank 0 over NCCL asynchronously behind the next micro-batch and streamed to pinned host memory
        (SURVEY.md 8e)."""
        if world == 1:
            out = None
            for _ in range(k):
                out = step(*load(0, B))
            return out
        res = clip.segment_clip(sam.segment_bits, load, world * B * k, RESIZE, ORIGINAL, micro_batch=B,
                                max_prompts_per_frame=NP, threshold=0.0, device=dev, batch_sink=batch_sink,
                                keep=False, prefetch=prefetch, checksum_fn=sam.checksum64)
        clip_stats["masks_checked"] += res.masks_checked
        clip_stats["checksum_mismatches"] += res.checksum_mismatches
        clip_stats["frames"] += res.frames_done
        return res

    def resident(lo, hi):
        return frames, segs

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run_steps(max(args.warmup, 0), resident)
    barrier()
    handle = sam._handle

    # ---- timed region 1: inputs resident in HBM, CUDA events, max over ranks
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = sam.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    run_steps(args.steps, resident)
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    launches = sam.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    fps = world * B * args.steps / (ms_total * 1e-3)

    # ---- timed region 2: end to end from pinned host buffers (H2D frames + prompts, D2H packed masks + heads)
    h2d = host_frames.numel() * host_frames.element_size() + host_seg.numel() * 4
    d2h = lay.nbytes

    def from_host(lo, hi):
        fr = host_frames.to(dev, non_blocking=True)
        sg_all = host_seg.to(dev, non_blocking=True)
        return fr, [sg_all[i * NP:(i + 1) * NP] for i in range(B)]

    first_block = {}

    def host_sink(src, lo, cnt, bits, heads):
        """N > 1, rank 0 (consumer thread): the gathered round already sits in pinned host memory; keep the first
        block of the last rank for the cross-GPU identity check below."""
        if src == world - 1 and "bits" not in first_block and lo == clip.shard_range(world * B * args.steps, src, world)[0]:
            first_block["bits"], first_block["heads"], first_block["cnt"] = bits.clone(), heads.clone(), list(cnt)

    # N = 1: the public host-to-host API (clip.HostPipeline): every step uploads ITS frames + prompts from pinned host
    # memory and downloads ITS message (packed masks + heads), the copies of neighbouring steps overlapping the kernels.
    # N > 1: segment_clip with host loads (prefetched on an upload stream) and the host-side consumer on rank 0.
    pipe = clip.HostPipeline(sam.segment_bits, B, NP, RESIZE, ORIGINAL, dev) if world == 1 else None

    def e2e_run(k, sink=None):
        if world == 1:
            for _ in range(k):
                pipe.submit(host_frames, host_seg)
            pipe.flush()                         # the caller consumes the masks on the host
        else:
            run_steps(k, from_host, sink, prefetch=True)

    e2e_run(2)
    barrier()
    t0 = time.perf_counter()
    e2e_run(args.steps, host_sink)
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_fps = world * B * args.steps / float(t.item())

    # ---- what rank 0 received is what the producers computed, and what a single GPU computes (untimed)
    gather_verified = None
    if world > 1 and rank == 0:
        src = world - 1
        fr = make_frames(torch, src * B, B).to(torch.bfloat16).to(dev)
        sg = [make_seg(torch, src * B + i, NP).to(dev) for i in range(B)]
        own = sam.segment_bits(fr, sg, RESIZE, ORIGINAL, threshold=0.0)
        same = ("bits" in first_block and first_block["cnt"] == [NP] * B
                and torch.equal(first_block["bits"][0], own["bits_left"].cpu())
                and torch.equal(first_block["bits"][1], own["bits_right"].cpu())
                and torch.equal(first_block["heads"][0], own["taxonomy"].cpu())
                and torch.equal(first_block["heads"][1], own["iou_left"].cpu())
                and torch.equal(first_block["heads"][2], own["iou_right"].cpu()))
        gather_verified = {
            "ok": bool(same and clip_stats["checksum_mismatches"] == 0 and clip_stats["masks_checked"] > 0),
            "frames_gathered": clip_stats["frames"], "masks_checksummed": clip_stats["masks_checked"],
            "checksum_mismatches": clip_stats["checksum_mismatches"],
            "cross_gpu": {"frames": B, "source_rank": src, "bit_identical_to_rank0_recompute": bool(same)},
            "how": "b200sam_checksum64 of every (side, prompt) mask on the producing rank vs. on the bytes the NCCL "
                   "gather delivered; plus rank 0 recomputes the last rank's first micro-batch on its own GPU"}

    # ---- B = 1 latency through the reference's own call sequence (LISA.evaluate, LISA.py:487-532)
    latency = None
    if rank == 0:
        latency = measure_latency_b1(torch, sam, frames[:1], segs[0])

    # ---- per-launch CUDA events (b200sam_profile_*) over the same steps -> roofline of the dominant kernel
    kernels, roofline = {}, None
    if rank == 0:
        lib.b200sam_profile_enable(handle, 1)
        psteps = min(args.steps, 3)
        for _ in range(psteps):
            step(frames, segs)
        torch.cuda.synchronize()
        prof = read_profile(lib, handle)
        lib.b200sam_profile_enable(handle, 0)
        fl = flops_per_frame(NP)
        peaks = load_peaks()
        tot_ms = sum(v[0] for v in prof.values()) / psteps
        for name, (ms, cnt) in prof.items():
            if cnt == 0:
                continue
            k = {"ms_per_step": ms / psteps, "launches_per_step": cnt // psteps, "share": (ms / psteps) / tot_ms}
            if name in fl:
                k["tflops"] = fl[name] * B / (ms / psteps * 1e-3) / 1e12
            kernels[name] = k
        dom = max((n for n in kernels if n in fl), key=lambda n: kernels[n]["ms_per_step"])
        kd = kernels[dom]
        roofline = {"bound": "tensor", "kernel": dom, "achieved": kd["tflops"], "peak": peaks["bf16_sustained"],
                    "unit": "TFLOP/s", "frac": kd["tflops"] / peaks["bf16_sustained"], "traffic": ncu_traffic(dom),
                    "peak_source": peaks["source"] + ", sustained figure (kernel timed inside a long step)",
                    "flops_per_launch": fl[dom] * B / kd["launches_per_step"],
                    "avg_launch_ms": kd["ms_per_step"] / kd["launches_per_step"],
                    "whole_step_tflops": fl["total"] * B / (ms_total / args.steps * 1e-3) / 1e12,
                    "whole_step_frac": fl["total"] * B / (ms_total / args.steps * 1e-3) / 1e12 / peaks["bf16_sustained"],
                    "whole_step_frac_of": {
                        "measured_sustained": fl["total"] * B / (ms_total / args.steps * 1e-3) / 1e12 / peaks["bf16_sustained"],
                        "measured_burst": fl["total"] * B / (ms_total / args.steps * 1e-3) / 1e12 / peaks["bf16_burst"],
                        "nominal_2250": fl["total"] * B / (ms_total / args.steps * 1e-3) / 1e12 / 2250.0}}

    aux = None
    if rank == 0 and world == 1:
        out_u8 = sam.segment(frames, segs, resize, original, threshold=0.0)
        aux = measure_callers(torch, sam, dev, out_u8, B, NP)
        del out_u8

    sd_cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sd_cpu = {k: v.detach().float().cpu() for k, v in sam.state_dict().items()}

    # ---- the trainable part of the branch (both decoders forward + backward, postprocess, mask
    # losses) for the same 8 frames x 3 prompts; an extra key, not part of the headline metric
    train = None
    if rank == 0 and world == 1 and not args.no_train:
        import types
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        bench_train = __import__("bench_train")
        del sam
        torch.cuda.empty_cache()
        train = bench_train.run(types.SimpleNamespace(
            frames=B, prompts=NP, steps=3, warmup=2, original=list(original[0]), no_eager=args.no_cpu_baseline,
            precision="tf32", cpu_sample=0 if args.no_cpu_baseline else 1))
        sam = None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_block(torch, sd_cpu, NP)

    if rank == 0:
        line = {
            "metric": "frames/s ViT-H encode + bimanual mask decode @1024^2", "value": fps, "unit": "frames/s",
            "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(args, B),
            "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "kernels": kernels, "callers": aux, "latency_b1_ms": latency, "gather_verified": gather_verified,
            "train_step": train,
        }
        emit_line(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


_REAL_STDOUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL's version banner at communicator
    creation, warnings of native extensions), so file descriptor 1 is pointed at stderr for the whole run and the
    result line is written to the saved descriptor by emit_line()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit_line(text: str) -> None:
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        print(text, flush=True)
    else:
        os.write(_REAL_STDOUT, (text + "\n").encode())


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=8, help="frames per step per GPU (configs[1]: 8)")
    ap.add_argument("--prompts", type=int, default=3, help="[SEG] prompts per frame")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the decoder-training timing (extra key train_step)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    args.warmup = max(args.warmup, 3)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
