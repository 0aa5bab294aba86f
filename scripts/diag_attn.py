"""Diagnostics for the global attention op: structured inputs that isolate the P.V accumulation, the normalisation and
the bias paths.  Run on a GPU box: python scripts/diag_attn.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import gpu_util as G

def run(qkv, rel_h, rel_w, nb, S, heads, hd):
    out = torch.empty((nb * S * S, heads * hd), dtype=torch.bfloat16, device="cuda")
    rc = G.lib().b200sam_op_attention(G.P(qkv), nb, S, heads, hd, G.P(rel_h), G.P(rel_w), G.P(out), G.stream())
    torch.cuda.synchronize()
    assert rc == 0
    return out.float()

for hd in (64, 80):
    nb, S, heads = 1, 64, 1
    T = S * S
    g = torch.Generator(device="cpu").manual_seed(1)
    zero_rel = torch.zeros((2 * S - 1, hd), device="cuda")
    # (a) q = 0, V = 1 -> out = 1
    qkv = torch.zeros((T, 3 * hd), dtype=torch.bfloat16, device="cuda")
    qkv[:, 2 * hd:] = 1.0
    o = run(qkv, zero_rel, zero_rel, nb, S, heads, hd)
    print(hd, "V=1:", o.min().item(), o.max().item())
    # (b) q = 0, V[k, c] = key row index / 64 -> out = mean = 31.5/64
    qkv = torch.zeros((T, 3 * hd), dtype=torch.bfloat16, device="cuda")
    ky = (torch.arange(T, device="cuda") // 64).float() / 64
    qkv[:, 2 * hd:] = ky[:, None].bfloat16()
    o = run(qkv, zero_rel, zero_rel, nb, S, heads, hd)
    print(hd, "V=ky/64 (expect %.4f):" % ky.mean().item(), o.min().item(), o.max().item())
    # (c) V[k, c] = c / hd -> out[:, c] = c / hd (column mapping of the P.V product)
    qkv = torch.zeros((T, 3 * hd), dtype=torch.bfloat16, device="cuda")
    qkv[:, 2 * hd:] = (torch.arange(hd, device="cuda").float() / hd).bfloat16()[None, :]
    o = run(qkv, zero_rel, zero_rel, nb, S, heads, hd)
    ref = (torch.arange(hd, device="cuda").float() / hd).bfloat16().float()
    print(hd, "V=c/hd: max err", (o - ref[None]).abs().max().item(), "worst col", (o - ref[None]).abs().max(0).values.argmax().item())
    # (d) random q,k, V = 1 -> out = 1 (normalisation consistent with the accumulated P)
    qkv = (torch.randn((T, 3 * hd), generator=g) * 1.5).bfloat16().cuda()
    qkv[:, 2 * hd:] = 1.0
    o = run(qkv, zero_rel, zero_rel, nb, S, heads, hd)
    print(hd, "rand qk, V=1:", o.min().item(), o.max().item())
    # (e) random q, k = 0, rel_w only, V = kx/64: tests the w-bias path alone
    qkv = (torch.randn((T, 3 * hd), generator=g) * 1.5).bfloat16().cuda()
    qkv[:, hd:2 * hd] = 0
    kx = (torch.arange(T, device="cuda") % 64).float() / 64
    qkv[:, 2 * hd:] = kx[:, None].bfloat16()
    rel_w = (torch.randn((2 * S - 1, hd), generator=g) * 0.3).cuda()
    o = run(qkv, zero_rel, rel_w, nb, S, heads, hd)
    q = qkv[:, :hd].float()
    Rw = rel_w.bfloat16().float()
    qx = torch.arange(T, device="cuda") % 64
    idx = qx[:, None] - torch.arange(64, device="cuda")[None, :] + 63
    bw = torch.einsum("tc,tkc->tk", q, Rw[idx])
    p = torch.softmax(bw, -1)
    ref = (p * (torch.arange(64, device="cuda").float() / 64).bfloat16().float()[None]).sum(-1)
    print(hd, "w-bias only: max err", (o[:, 0] - ref).abs().max().item())
