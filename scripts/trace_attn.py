#!/usr/bin/env python
"""Phase timeline of CTA 0 of the windowed attention kernel (tracing variant, include/b200sam_debug.h): clock64() at
the phase boundaries of the first 16 problems, per warp.  Prints, per problem, the cycles each softmax warp spent in
each phase and when the MMA warp issued S / P.V.  Usage: python scripts/trace_attn.py [--variant 2] [--hd 80]"""
import argparse
import ctypes as C
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

_lib = importlib.import_module("2handedafforder_b200._lib")
ITS, TAGS = 16, 10


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", type=int, default=2)
    ap.add_argument("--hd", type=int, default=80)
    ap.add_argument("--heads", type=int, default=16)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--S", type=int, default=14)
    a = ap.parse_args()
    lib = _lib.load()
    S, nb = a.S, a.batch * (25 if a.S == 14 else 1)
    g = torch.Generator().manual_seed(30 + S)
    rows = nb * S * S
    qkv = (torch.randn((rows, 3 * a.heads * a.hd), generator=g) * 1.5).to(torch.bfloat16).cuda()
    rel_h = (torch.randn((2 * S - 1, a.hd), generator=g) * 0.3).cuda()
    rel_w = (torch.randn((2 * S - 1, a.hd), generator=g) * 0.3).cuda()
    out = torch.zeros((rows, a.heads * a.hd), dtype=torch.bfloat16, device="cuda")
    trace = torch.zeros((16, ITS, TAGS), dtype=torch.int64, device="cuda")
    assert lib.b200sam_debug_set_trace(C.c_void_p(trace.data_ptr())) == 0
    assert lib.b200sam_debug_set_variant(0 if S == 14 else 1, a.variant) == 0
    ms = C.c_float()
    rc = lib.b200sam_debug_attention_bench(C.c_void_p(qkv.data_ptr()), nb, S, a.heads, a.hd,
                                           C.c_void_p(rel_h.data_ptr()), C.c_void_p(rel_w.data_ptr()),
                                           C.c_void_p(out.data_ptr()), 1, C.byref(ms),
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    lib.b200sam_debug_set_variant(0 if S == 14 else 1, 0)
    lib.b200sam_debug_set_trace(None)
    assert rc == 0
    t = trace.cpu().numpy()
    t0 = int(t[t > 0].min())
    rel = lambda v: int(v - t0) if v > 0 else -1
    print(f"ms per launch (traced variant) {ms.value:.4f}")
    print("TMA warp: it: qk issue, v issue")
    for it in range(ITS):
        print(f"  it{it:2d}: qk {rel(t[0, it, 0]):7d}  v {rel(t[0, it, 1]):7d}   landed: qk {rel(t[2, it, 0]):7d} "
              f"(+{rel(t[2, it, 0]) - rel(t[0, it, 0])})  v {rel(t[2, it, 1]):7d} (+{rel(t[2, it, 1]) - rel(t[0, it, 1])})")
    print("MMA warps (unit 0 | unit 1): qk_full seen, o_free seen, S issued, G issued, p_full seen, PV issued")
    for it in range(ITS):
        print(f"  it{it:2d}: " + " ".join(f"{rel(t[1, it, k]):7d}" for k in (2, 3, 4, 0, 5, 1)) + " | " +
              " ".join(f"{rel(t[3, it, k]):7d}" for k in (2, 3, 4, 0, 5, 1)))
    print("observer: s_full[0], s_full[1], o_full[0], o_full[1] completion times")
    for it in range(ITS):
        print(f"  it{it:2d}: " + " ".join(f"{rel(t[2, it, k]):7d}" for k in (2, 3, 4, 5)))
    print("epilogue warps 12-15: o_full(0) seen, o_free(0) arrive, o_full(1) seen, o_free(1) arrive")
    for it in range(2, 8):
        print(f"  it{it:2d}: " + "  ".join("w%d: %s" % (w, "/".join(str(rel(t[w, it, k])) for k in range(4)))
                                          for w in (12, 13, 14, 15)))
    names = ["g_full", "gather", "s_wait", "pass1+M", "pass2", "p_arrive"]
    print("softmax warps: per problem, time of each mark and deltas")
    for w in (4, 8, 10):
        print(f" warp {w} (unit {(w - 4) >> 2})")
        for it in range(2, 10):
            m = [rel(t[w, it, k]) for k in range(6)]
            d = [m[k] - m[k - 1] for k in range(1, 6)]
            nxt = rel(t[w, it + 1, 0]) - m[5] if it + 1 < ITS else -1
            print(f"  it{it:2d} g_full@{m[0]:7d}  d: " + " ".join(f"{n}={v}" for n, v in zip(names[1:], d)) +
                  f"  wait_next_G={nxt}  period={rel(t[w, it + 1, 0]) - m[0] if it + 1 < ITS else -1}")
    # average period over its 2..14
    for w in (4, 8):
        per = [(t[w, it + 1, 0] - t[w, it, 0]) for it in range(2, ITS - 1)]
        print(f"warp {w}: mean period {sum(per) / len(per):.0f} cycles")
    print(json.dumps({"ms": ms.value}))


if __name__ == "__main__":
    main()
