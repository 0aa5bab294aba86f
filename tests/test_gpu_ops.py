"""Operator-level parity of the sm_100a kernels (through the C ABI) against plain fp32 PyTorch on the GPU.

Tolerances: bf16-operand GEMMs with fp32 accumulation are compared with an fp32 matmul of the SAME bf16-rounded
operands (so only accumulation order and output rounding differ): 4e-3 of the output range for bf16 outputs
(2^-8 output rounding), 2e-4 for fp32 outputs.  TF32 GEMMs: 2e-3 of range against full fp32.
"""
import math

import pytest
import torch

from oracle import synth
from tests import gpu_util as G

pytestmark = pytest.mark.gpu


def _rand(shape, seed, scale=1.0, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype).cuda()


GEMM_SHAPES = [(128, 256, 64), (300, 960, 320), (4096, 3840, 1280), (4100, 1280, 5120), (4096, 256, 2304)]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
@pytest.mark.parametrize("mode", [0, 1])
def test_gemm_bf16_out(M, N, K, mode):
    A = _rand((M, K), 1, 1.0, torch.bfloat16)
    W = _rand((N, K), 2, 1.0 / math.sqrt(K), torch.bfloat16)
    bias = _rand((N,), 3, 0.5)
    ref = A.float() @ W.float().t() + bias
    if mode == 1:
        ref = torch.nn.functional.gelu(ref)
    out = G.op_gemm(A, W, mode, bias=bias)
    assert G.rel_err(out, ref) < 4e-3
    simt = G.op_gemm(A[:256], W, mode, bias=bias, simt=True)
    assert G.rel_err(simt, ref[:256]) < 4e-3


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_f32_residual_in_place(M, N, K):
    A = _rand((M, K), 4, 1.0, torch.bfloat16)
    W = _rand((N, K), 5, 1.0 / math.sqrt(K), torch.bfloat16)
    bias = _rand((N,), 6, 0.5)
    X = _rand((M, N), 7)
    ref = X + A.float() @ W.float().t() + bias
    out = X.clone()
    G.op_gemm(A, W, 2, bias=bias, addend=out, out=out)      # x = x + proj(...) in place
    assert G.rel_err(out, ref) < 2e-4


@pytest.mark.parametrize("M,N,K,add_mod", [
    (4096, 1280, 256, 0),       # ViT-H width: compile-time strides, every warp full
    (1100, 1280, 128, 0),       # last pair tile holds 76 rows: full, partial (12 rows) and empty warps
    (1100, 768, 64, 1024),      # ViT-B width, broadcast addend wrapping inside the last tile
    (2048, 1024, 64, 0),        # ViT-L width
    (1100, 512, 64, 0),         # any other width: run-time strides
    (300, 256, 64, 0),          # below 1024 rows: the single-CTA kernel
])
def test_gemm_fold_producer_epilogue(M, N, K, add_mod):
    """Residual-stream epilogue with the producer half of the folded LayerNorm: fp32 row, its bf16
    copy and the per-row partial {sum, sum of squares} of every 128-column slice, through b200sam_op_gemm_fold_producer.
    Tolerance: bf16 operands with fp32 accumulation against fp32 torch, 2e-4 of the largest value (as the GM_F32 tests);
    the bf16 copy must be the round-to-nearest-even of the stored fp32 row exactly; the sums within 1e-5 relative of the
    fp64 sums of the stored row (different summation order only)."""
    A = _rand((M, K), 31, 1.0, torch.bfloat16)
    W = _rand((N, K), 32, 1.0 / math.sqrt(K), torch.bfloat16)
    bias = _rand((N,), 33, 0.5)
    if add_mod:
        add = _rand((add_mod, N), 34)
        ref = A.float() @ W.float().t() + bias + add[torch.arange(M, device=add.device) % add_mod]
        out = torch.full((M, N), float("nan"), dtype=torch.float32, device="cuda")
        addend = add
    else:
        X = _rand((M, N), 34)
        ref = X + A.float() @ W.float().t() + bias
        out = X.clone()          # in place, as proj / lin2 update the residual stream
        addend = out
    slots = 2 * ((N + 255) // 256)
    pad = 7                      # rows behind M must stay untouched
    xb = torch.full((M + pad, N), float("nan"), dtype=torch.bfloat16, device="cuda")
    stats = torch.full((M + pad, slots, 2), float("nan"), dtype=torch.float32, device="cuda")
    rc = G.lib().b200sam_op_gemm_fold_producer(0, G.P(A), A.stride(0), G.P(W), W.stride(0), M, N, K, G.P(bias),
                                               G.P(addend), addend.stride(0), add_mod, G.P(out), out.stride(0),
                                               G.P(xb), G.P(stats), G.stream())
    assert rc == 0
    torch.cuda.synchronize()
    assert G.rel_err(out, ref) < 2e-4
    assert torch.equal(xb[:M], out.bfloat16())
    assert torch.isnan(xb[M:].float()).all() and torch.isnan(stats[M:]).all()
    o64 = out.double()
    for j in range(slots):
        sl = o64[:, j * 128:(j + 1) * 128]           # every N here is a multiple of 256: one 128-column slice per slot
        s1, s2 = sl.sum(1), (sl * sl).sum(1)
        assert float((stats[:M, j, 0].double() - s1).abs().max()) <= 1e-5 * float(sl.abs().sum(1).max()) + 1e-12
        assert float((stats[:M, j, 1].double() - s2).abs().max()) <= 1e-5 * float(s2.max()) + 1e-12


def test_gemm_fold_producer_rejects_what_the_epilogue_does_not_handle():
    A = _rand((1024, 64), 35, 1.0, torch.bfloat16)
    W = _rand((256, 64), 36, 0.1, torch.bfloat16)
    out = torch.zeros((1024, 256), dtype=torch.float32, device="cuda")
    xb = torch.zeros((1024, 256), dtype=torch.bfloat16, device="cuda")
    stats = torch.zeros((1024, 2, 2), dtype=torch.float32, device="cuda")
    call = lambda addend, add_mod: G.lib().b200sam_op_gemm_fold_producer(
        0, G.P(A), 64, G.P(W), 64, 1024, 256, 64, None, G.P(addend), 256, add_mod, G.P(out), 256, G.P(xb), G.P(stats),
        G.stream())
    assert call(None, 0) == -1            # B200SAM_ERR_INVALID: the addend is required
    assert call(out, 48) == -1            # add_mod must be a multiple of 32
    assert call(out, 0) == 0
    torch.cuda.synchronize()


@pytest.mark.parametrize("M,N,K,gelu", [
    (4096, 3840, 1280, False),   # ViT-H qkv
    (4096, 1024, 256, True),     # lin1-like with GELU
    (1100, 512, 256, True),      # last pair tile: full, partial and empty warps
    (1100, 512, 384, False),     # K % 256 = 128: the last stats slot is the producer's zero slot
    (300, 256, 128, False),      # below 1024 rows: the single-CTA kernel
])
def test_gemm_fold_consumer_epilogue(M, N, K, gelu):
    """qkv / lin1 with the LayerNorm in front of them folded in, through b200sam_op_gemm_fold_consumer:
    weight folding kernel + GEMM + epilogue.  Two checks.  (1) Against the definition LayerNorm(x) W^T + b [GELU] in
    fp32 torch: 1.5e-2 of the largest value, the bound the path tests use for one block in bf16 (x and gamma.W are
    rounded to bf16).  (2) Against the same algebra evaluated in fp64 from the SAME rounded operands and sums, which
    leaves only the fp32 accumulation order and the bf16 rounding of the result: 2**-8 of the largest value."""
    eps = 1e-6
    x = _rand((M, K), 41) + 0.3
    W = _rand((N, K), 42, 1.0 / math.sqrt(K))
    gamma, beta, bias = 1.0 + _rand((K,), 43, 0.2), _rand((K,), 44, 0.2), _rand((N,), 45, 0.5)
    xb = x.bfloat16()
    slots = 2 * ((K + 255) // 256)
    stats = torch.zeros((M, slots, 2), dtype=torch.float32, device="cuda")
    for j in range((K + 127) // 128):
        sl = x[:, j * 128:(j + 1) * 128]
        stats[:, j, 0], stats[:, j, 1] = sl.sum(1), (sl * sl).sum(1)
    wf = torch.empty((N, K), dtype=torch.bfloat16, device="cuda")
    cd = torch.empty((2 * N,), dtype=torch.float32, device="cuda")
    pad = 5
    out = torch.full((M + pad, N), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = G.lib().b200sam_op_gemm_fold_consumer(0, int(gelu), G.P(xb), K, G.P(W), G.P(gamma), G.P(beta), G.P(bias), eps,
                                               G.P(stats), M, N, K, G.P(wf), G.P(cd), G.P(out), N, G.stream())
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.isnan(out[M:].float()).all()
    # the folded weights
    assert torch.equal(wf, (W * gamma).bfloat16())
    assert G.rel_err(cd[:N], wf.double().sum(1)) < 1e-5 * math.sqrt(K)
    assert G.rel_err(cd[N:], W.double() @ beta.double() + bias.double()) < 1e-5
    act = torch.nn.functional.gelu if gelu else (lambda t: t)
    ref = act(torch.nn.functional.layer_norm(x, (K,), gamma, beta, eps) @ W.t() + bias)
    assert G.rel_err(out[:M], ref) < 1.5e-2
    s1, s2 = stats[:, :, 0].double().sum(1), stats[:, :, 1].double().sum(1)
    mu = s1 / K
    rstd = 1.0 / torch.sqrt((s2 / K - mu * mu).clamp_min(0) + eps)
    emul = rstd[:, None] * (xb.double() @ wf.double().t()) - (rstd * mu)[:, None] * cd[:N].double() + cd[N:].double()
    assert G.rel_err(out[:M], act(emul)) < 2.0 ** -8


def test_gemm_f32_broadcast_addend():
    """patch-embed epilogue: + bias + pos_embed[row % 4096] (image_encoder.py:111-113)."""
    M, N, K = 2 * 4096, 320, 768
    A = _rand((M, K), 8, 1.0, torch.bfloat16)
    W = _rand((N, K), 9, 1.0 / math.sqrt(K), torch.bfloat16)
    bias = _rand((N,), 10, 0.5)
    pos = _rand((4096, N), 11)
    ref = A.float() @ W.float().t() + bias + pos.repeat(2, 1)
    out = G.op_gemm(A, W, 2, bias=bias, addend=pos, add_mod=4096)
    assert G.rel_err(out, ref) < 2e-4


def test_gemm_window_unpartition_rowmap():
    """proj epilogue scatters window-ordered rows back to token order, dropping pad rows
    (window_unpartition, image_encoder.py:291-318)."""
    B, D = 2, 256
    A = _rand((B * 4900, D), 12, 1.0, torch.bfloat16)
    W = _rand((D, D), 13, 1.0 / math.sqrt(D), torch.bfloat16)
    X = _rand((B * 4096, D), 14)
    full = (A.float() @ W.float().t()).reshape(B, 5, 5, 14, 14, D).permute(0, 1, 3, 2, 4, 5).reshape(B, 70, 70, D)
    ref = X + full[:, :64, :64].reshape(B * 4096, D)
    out = X.clone()
    G.op_gemm(A, W, 2, addend=out, out=out, rowmap=1)
    assert G.rel_err(out, ref) < 2e-4


@pytest.mark.parametrize("tf32", [False, True])
def test_gemm_layernorm_epilogue(tf32):
    M, N, K = 4096 + 37, 256, 128 if tf32 else 1280
    dt = torch.float32 if tf32 else torch.bfloat16
    A = _rand((M, K), 15, 1.0, dt)
    W = _rand((N, K), 16, 1.0 / math.sqrt(K), dt)
    bias = _rand((N,), 17, 0.5)
    add = _rand((M, N), 18)
    g, b = 1 + 0.1 * _rand((N,), 19), 0.05 * _rand((N,), 20)
    ref = torch.nn.functional.layer_norm(add + A.float() @ W.float().t() + bias, (N,), g, b, 1e-5)
    out = G.op_gemm(A, W, 3, tf32=tf32, bias=bias, addend=add, ln_g=g, ln_b=b, eps=1e-5)
    assert G.rel_err(out, ref) < (2e-3 if tf32 else 3e-4)


@pytest.mark.parametrize("M,N,K", [(4096, 256, 256), (5000, 128, 256), (333, 256, 128)])
def test_gemm_tf32(M, N, K):
    A = _rand((M, K), 21)
    W = _rand((N, K), 22, 1.0 / math.sqrt(K))
    tab = _rand((4096, N), 23)
    ref = A @ W.t() + tab[torch.arange(M).cuda() % 4096]
    out = G.op_gemm(A, W, 2, tf32=True, addend=tab, add_mod=4096)
    assert G.rel_err(out, ref) < 2e-3
    exact = G.op_gemm(A[:128], W, 2, tf32=True, addend=tab, add_mod=4096, simt=True)
    assert G.rel_err(exact, ref[:128]) < 1e-5


def test_gemm_rejects_bad_shapes():
    A = _rand((128, 60), 24, 1.0, torch.bfloat16)
    W = _rand((256, 60), 25, 1.0, torch.bfloat16)
    out = torch.empty((128, 256), dtype=torch.bfloat16, device="cuda")
    rc = G.lib().b200sam_op_gemm(0, 0, 0, G.P(A), 60, G.P(W), 60, 128, 256, 60, None, None, 0, 0, None, None, 0.0,
                                 G.P(out), 256, 0, 0, G.stream())
    assert rc == -1   # K % 8 != 0 is an error, never a fallback


@pytest.mark.parametrize("D", [256, 320, 768, 1280])
@pytest.mark.parametrize("windowed", [0, 1])
def test_layernorm(D, windowed):
    B = 2
    x = _rand((B * 4096, D), 26, 2.0) + 0.3
    g, b = 1 + 0.1 * _rand((D,), 27), 0.05 * _rand((D,), 28)
    rows_out = B * 4900 if windowed else B * 4096
    out = torch.zeros((rows_out, D), dtype=torch.bfloat16, device="cuda")
    rc = G.lib().b200sam_op_layernorm(G.P(x), B * 4096, D, G.P(g), G.P(b), 1e-6, G.P(out), windowed, G.stream())
    assert rc == 0
    torch.cuda.synchronize()
    ref = torch.nn.functional.layer_norm(x, (D,), g, b, 1e-6)
    if windowed:   # window_partition with zero pad AFTER the norm (image_encoder.py:179-183,263-288)
        pad = torch.zeros((B, 70, 70, D), device="cuda")
        pad[:, :64, :64] = ref.reshape(B, 64, 64, D)
        ref = pad.reshape(B, 5, 14, 5, 14, D).permute(0, 1, 3, 2, 4, 5).reshape(B * 4900, D)
    assert G.rel_err(out, ref) < 4e-3
    if windowed:
        assert float(out.float().reshape(B, 25, 196, D)[:, 24, 195].abs().max()) == 0.0   # pad rows untouched


def _attention_ref(qkv, nb, S, heads, hd, rel_h, rel_w):
    """Attention.forward with add_decomposed_rel_pos (image_encoder.py:235-260,354-392) in fp32."""
    T = S * S
    x = qkv.float().reshape(nb, T, 3, heads, hd).permute(2, 0, 3, 1, 4)
    q, k, v = x[0], x[1], x[2]
    attn = (q * hd ** -0.5) @ k.transpose(-2, -1)
    idx = (torch.arange(S)[:, None] - torch.arange(S)[None, :] + S - 1).cuda()
    Rh = rel_h.bfloat16().float()[idx]
    Rw = rel_w.bfloat16().float()[idx]
    rq = q.reshape(nb, heads, S, S, hd)
    bh = torch.einsum("bnhwc,hkc->bnhwk", rq, Rh)
    bw = torch.einsum("bnhwc,wkc->bnhwk", rq, Rw)
    attn = attn.reshape(nb, heads, S, S, S, S) + bh[..., :, None] + bw[..., None, :]
    attn = torch.softmax(attn.reshape(nb, heads, T, T), -1)
    return (attn @ v).permute(0, 2, 1, 3).reshape(nb * T, heads * hd)


@pytest.mark.parametrize("S,nb,heads,hd", [(14, 50, 4, 64), (14, 27, 3, 80), (64, 2, 2, 64), (64, 1, 3, 80)])
def test_attention_relpos(S, nb, heads, hd):
    rows = nb * S * S
    qkv = _rand((rows, 3 * heads * hd), 30, 1.5, torch.bfloat16)
    rel_h = _rand((2 * S - 1, hd), 31, 0.3)
    rel_w = _rand((2 * S - 1, hd), 32, 0.3)
    out = torch.empty((rows, heads * hd), dtype=torch.bfloat16, device="cuda")
    rc = G.lib().b200sam_op_attention(G.P(qkv), nb, S, heads, hd, G.P(rel_h), G.P(rel_w), G.P(out), G.stream())
    assert rc == 0
    torch.cuda.synchronize()
    ref = _attention_ref(qkv, nb, S, heads, hd, rel_h, rel_w)
    # P is fed to the PV product as bf16 and the output is stored as bf16
    assert G.rel_err(out, ref) < 1e-2
    assert float((out.float() - ref).abs().mean() / ref.abs().mean()) < 5e-3


@pytest.mark.parametrize("S,nb,heads,hd", [(64, 1, 2, 80), (14, 9, 2, 64)])
def test_attention_relpos_checkpoint_like_magnitudes(S, nb, heads, hd):
    """A checkpoint-like range instead of the N(0, 0.02) tables of random init: q of sigma 4 (extremes ~16) against
    rel-pos entries of sigma 0.5 gives biases of sigma ~16, extremes ~70.  The global blocks keep q.Rh / q.Rw as fp16
    (the w part is a tensor-core operand): documented bound |d exponent| <= 2^-11 (|bh| + |bw|), i.e. <= ~0.07 here -
    still 4x tighter than the reference's own bf16 mode, which rounds the same projections to bf16 (2^-9).  The
    windowed kernel reads them from TMEM in fp32.  Measured on B200: sigma-12 q (biases to ~400) gives 0.19 absolute
    output error in the global kernel, which is why the bound is stated rather than hidden."""
    rows = nb * S * S
    qkv = _rand((rows, 3 * heads * hd), 40, 1.0, torch.bfloat16)
    qkv[:, : heads * hd] *= 4.0
    rel_h = _rand((2 * S - 1, hd), 41, 0.5)
    rel_w = _rand((2 * S - 1, hd), 42, 0.5)
    out = torch.empty((rows, heads * hd), dtype=torch.bfloat16, device="cuda")
    rc = G.lib().b200sam_op_attention(G.P(qkv), nb, S, heads, hd, G.P(rel_h), G.P(rel_w), G.P(out), G.stream())
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    ref = _attention_ref(qkv, nb, S, heads, hd, rel_h, rel_w)
    # near one-hot softmax rows: compare on an absolute scale (|v| ~ 1)
    assert float((out.float() - ref).abs().max()) < 6e-2
    assert float((out.float() - ref).abs().mean() / ref.abs().mean()) < 1e-2


@pytest.mark.parametrize("arch_name,batch", [("tiny64", 1), ("tiny64", 3), ("tiny80", 2)])
def test_tma_gathered_convolutions_equal_the_im2col_path(arch_name, batch):
    """PatchEmbed Conv2d(k16,s16) and the neck's Conv2d 3x3 (image_encoder.py:395-426, 100-106) read their A tiles
    through multi-dimensional TMA tensor maps (gemm_tc.cuh AM_PATCH16 / AM_CONV3).  Same K order, same tiles, same
    accumulation order as the im2col copy + plain GEMM (debug variant 3 = 1), so the embeddings are bit-identical.
    tiny80 (width 320, not a multiple of 256) takes the im2col patch path but the gathered 3x3; fp32 frames take
    im2col for the patch embedding by design (the index map is also the cast)."""
    sam, arch, sd = G.build_sam_from_synth(arch_name, max_batch=batch, max_prompts=1)
    frames = synth.make_frames(3, batch).cuda()
    try:
        for dt in (torch.bfloat16, torch.float32):
            x = frames.to(dt)
            assert G.lib().b200sam_debug_set_variant(3, 1) == 0
            want = sam.get_visual_embs(x)
            assert G.lib().b200sam_debug_set_variant(3, 0) == 0
            got = sam.get_visual_embs(x)
            got2 = sam.get_visual_embs(x)           # second / third call: the CUDA-graph replay of the body
            got3 = sam.get_visual_embs(x)
            torch.cuda.synchronize()
            assert torch.isfinite(want.float()).all() and want.float().abs().max() > 0
            assert torch.equal(got, want) and torch.equal(got2, want) and torch.equal(got3, want)
    finally:
        G.lib().b200sam_debug_set_variant(3, 0)


@pytest.mark.parametrize("arch_name,batch", [("tiny64", 2), ("tiny80", 1)])
def test_layernorm_folded_into_qkv_and_lin1(arch_name, batch):
    """Shipped path (debug variant 5 = 0; 1 = stand-alone LayerNorm kernel): norm1 / norm2 (image_encoder.py:179,191) are
    folded into the GEMMs that consume them,
    LN(x) W^T + b = rstd * (x (gamma . W)^T) - rstd * mean * c + d, with the row sums produced by the residual-stream
    epilogues (gemm_tc.cuh GemmEpi::stats_out / stats_in).  Same mathematics, different rounding points (x instead of
    LN(x) is rounded to bf16), so the embeddings agree to the accuracy either path has against the fp32 oracle."""
    from oracle import sam_oracle as O
    sam, arch, sd = G.build_sam_from_synth(arch_name, max_batch=batch, max_prompts=1)
    frames = synth.make_frames(5, batch)
    ref = O.encode(sd, arch, frames, O.FP32)
    x = frames.cuda().to(torch.bfloat16)
    try:
        assert G.lib().b200sam_debug_set_variant(5, 1) == 0
        base = sam.get_visual_embs(x).float().cpu()
        assert G.lib().b200sam_debug_set_variant(5, 0) == 0
        fold = sam.get_visual_embs(x).float().cpu()
        fold2 = sam.get_visual_embs(x).float().cpu()          # graph replay
        fold3 = sam.get_visual_embs(x).float().cpu()
    finally:
        G.lib().b200sam_debug_set_variant(5, 0)
    ref_b = O.encode(sd, arch, frames.to(torch.bfloat16).float(), O.FP32)
    e_base, e_fold = O.logit_rel_err(base, ref_b), O.logit_rel_err(fold, ref_b)
    assert torch.equal(fold, fold2) and torch.equal(fold, fold3)
    assert e_base < 1.5e-2 and e_fold < 1.5e-2 and e_fold < 1.5 * e_base + 1e-3, (e_base, e_fold)
    assert O.logit_rel_err(fold, base) < 1.5e-2
    del ref


def test_pad_rows_written_by_the_previous_lin2_launch_give_identical_embeddings():
    """The 804 window-pad rows of a windowed block's qkv buffer (zero after norm1, so qkv = bias there;
    image_encoder.py:179-183) are written by the idle warp of the PREVIOUS block's lin2 GEMM (default) instead of a fill
    kernel behind every windowed qkv GEMM (debug variant 7 = 1).  Same bytes at the same place before the attention
    reads them: the embeddings are bit-identical, eager and as a CUDA graph.  Three blocks of width 256 so that block
    1 is a windowed block behind a lin2 launch on the CTA-pair kernel (the tiny golden shapes have none)."""
    arch = synth.Arch("tiny64x3", 256, 3, 4, (2,), 3)
    sam = G.pkg()._build_sam(arch.dim, arch.depth, arch.heads, list(arch.global_idx), max_batch=2, max_prompts=1)
    missing, unexpected = sam.load_state_dict(synth.make_state_dict(arch, seed=0), strict=True)
    assert not missing and not unexpected
    sam = sam.cuda()
    x = synth.make_frames(12, 2).cuda().to(torch.bfloat16)
    try:
        assert G.lib().b200sam_debug_set_variant(7, 1) == 0
        want = sam.get_visual_embs(x)
        assert G.lib().b200sam_debug_set_variant(7, 0) == 0
        got = [sam.get_visual_embs(x) for _ in range(4)]          # eager, capture, two replays
        torch.cuda.synchronize()
    finally:
        G.lib().b200sam_debug_set_variant(7, 0)
    assert all(torch.equal(g, want) for g in got)
    assert torch.isfinite(want.float()).all()


def test_programmatic_dependent_launch_gives_identical_embeddings():
    """Debug variant 6 = 1 launches the GEMMs, both attention kernels and the pad-row fill with the programmatic
    dependent launch attribute (their barrier / TMEM set-up overlaps the previous kernel's tail; griddepcontrol.wait
    before the first global access).  Ordering only: the embeddings are bit-identical, eager and as a CUDA graph."""
    sam, arch, sd = G.build_sam_from_synth("tiny80", max_batch=2, max_prompts=1)
    x = synth.make_frames(11, 2).cuda().to(torch.bfloat16)
    try:
        assert G.lib().b200sam_debug_set_variant(6, 0) == 0
        want = sam.get_visual_embs(x)
        assert G.lib().b200sam_debug_set_variant(6, 1) == 0
        got = [sam.get_visual_embs(x) for _ in range(4)]          # eager, capture, two replays
        torch.cuda.synchronize()
    finally:
        G.lib().b200sam_debug_set_variant(6, 0)
    assert all(torch.equal(g, want) for g in got)
