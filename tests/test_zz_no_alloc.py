"""b200sam_seg_project_noalloc (include/b200sam.h): the allocation-free variant of b200sam_seg_project, whose scratch is a
file-scope `static __device__` array instead of handle-owned cudaMalloc buffers.  It has no counterpart in the
reference; the check is against the existing entry point.

Collected last on purpose.  Called through the C ABI, two calls in a row with
different inputs (the array keeps the previous call's contents), compared with b200sam_seg_project on the same inputs
within the tolerance tests/test_gpu_callers.py uses for this operator (1e-4 of the largest reference value; both entry
points launch the same kernels on the same grid, so the results are in fact bit-identical)."""
import ctypes as C

import pytest
import torch

from tests import gpu_util as G

pytestmark = pytest.mark.gpu

SEG = 32003


def _inputs(seed, b, T, hdim, counts, dtype):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, 32000, (b, T), generator=g)
    for bi, n in enumerate(counts):
        ids[bi, torch.randperm(T - 1, generator=g)[:n] + 1] = SEG
    hidden = torch.randn((b, 255 + T - 1 + 2, hdim), generator=g).to(dtype)
    w1 = (torch.randn((hdim, hdim), generator=g) / hdim ** 0.5).to(dtype)
    b1 = torch.randn((hdim,), generator=g).to(dtype)
    w2 = (torch.randn((256, hdim), generator=g) / hdim ** 0.5).to(dtype)
    b2 = torch.randn((256,), generator=g).to(dtype)
    return [t.cuda().contiguous() for t in (hidden, ids, w1, b1, w2, b2)]


def _call(fn, h, hidden, ids, w1, b1, w2, b2, max_rows):
    b, L, hdim = hidden.shape
    T = ids.shape[1]
    dt = {torch.float32: 0, torch.bfloat16: 1}
    emb = torch.full((max_rows, 256), float("nan"), dtype=torch.float32, device="cuda")
    counts = torch.full((b + 1,), -1, dtype=torch.int32, device="cuda")
    rc = fn(h, G.P(hidden), dt[hidden.dtype], b, L, hdim, G.P(ids), T, SEG, G.P(w1), G.P(b1), G.P(w2), G.P(b2),
            dt[w1.dtype], G.P(emb), max_rows, G.P(counts), G.stream())
    return rc, emb, counts


def test_seg_project_noalloc_matches_seg_project_on_two_calls_in_a_row():
    sam = G.pkg()._build_sam(256, 2, 4, [1], max_batch=1, max_prompts=1).cuda()
    h = sam._ensure_handle_only(torch.device("cuda", torch.cuda.current_device()))
    lib = G.lib()
    # call 1: LLaVA-13B hidden size in bf16, 19 prompts over 4 images; call 2: smaller everything in fp32, so that stale
    # rows / counts / hidden activations of call 1 would show if anything depended on the array's previous contents
    cases = [(_inputs(5, 4, 64, 5120, (7, 0, 11, 1), torch.bfloat16), 24, [7, 0, 11, 1, 19]),
             (_inputs(6, 2, 40, 512, (2, 1), torch.float32), 8, [2, 1, 3])]
    outs = []
    for k, (inp, cap, want_counts) in enumerate(cases):
        if k == 1:
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
        rc, emb, counts = _call(lib.b200sam_seg_project_noalloc, h, *inp, cap)
        assert rc == 0, G.libmod().last_error(h)
        torch.cuda.synchronize()
        if k == 1:      # the set-up (cudaGetSymbolAddress, module load) happened in call 1: nothing may be allocated now
            assert torch.cuda.mem_get_info()[0] >= free_before
        outs.append((emb, counts))
        assert counts.tolist() == want_counts
    for (inp, cap, want_counts), (emb, counts) in zip(cases, outs):
        rc, ref_emb, ref_counts = _call(lib.b200sam_seg_project, h, *inp, cap)
        assert rc == 0, G.libmod().last_error(h)
        torch.cuda.synchronize()
        n = want_counts[-1]
        assert torch.equal(counts, ref_counts)
        assert torch.isfinite(ref_emb[:n]).all() and torch.isnan(emb[n:]).all()      # rows beyond the total untouched
        err = float((emb[:n] - ref_emb[:n]).abs().max())
        assert err <= 1e-4 * float(ref_emb[:n].abs().max()), err
        assert torch.equal(emb[:n], ref_emb[:n])                                      # same kernels, same grid
    # against the definition (fp32 torch on the GPU, TF32 off in tests/gpu_util.py) for the second case
    hidden, ids, w1, b1, w2, b2 = cases[1][0]
    rows = [hidden[bi, 255 + t - 1] for bi in range(ids.shape[0]) for t in range(1, ids.shape[1]) if ids[bi, t] == SEG]
    x = torch.stack(rows).float()
    want = torch.relu(x @ w1.float().t() + b1.float()) @ w2.float().t() + b2.float()
    assert float((outs[1][0][:3] - want).abs().max()) <= 1e-4 * float(want.abs().max())


def test_seg_project_noalloc_rejects_sizes_beyond_the_static_scratch():
    sam = G.pkg()._build_sam(256, 2, 4, [1], max_batch=1, max_prompts=1).cuda()
    h = sam._ensure_handle_only(torch.device("cuda", torch.cuda.current_device()))
    lib = G.lib()
    inp = _inputs(7, 1, 16, 512, (1,), torch.float32)
    rc, _, _ = _call(lib.b200sam_seg_project_noalloc, h, *inp, 65)            # max_rows > 64
    assert rc == -1 and "max_rows <= 64" in G.libmod().last_error(h)
    big = _inputs(8, 1, 16, 5128, (1,), torch.bfloat16)                        # hdim > 5120
    rc, _, _ = _call(lib.b200sam_seg_project_noalloc, h, *big, 8)
    assert rc == -1 and "hdim <= 5120" in G.libmod().last_error(h)
    rc, emb, counts = _call(lib.b200sam_seg_project_noalloc, h, *inp, 64)     # the limit itself is fine
    assert rc == 0 and counts.tolist() == [1, 1]
