"""Training forward / backward of the two mask decoders on libb200sam's fp32 primitives.

The reference trains `mask_decoder_left` / `mask_decoder_right` and `text_hidden_fcs` while the image encoder stays
frozen (2Haff/model/LISA.py:79-104); the SAM part of its training step is the per-image loop of LISA.py:284-325
(prompt encoder -> both decoders -> postprocess_masks, all under autograd) and the mask losses of LISA.py:16-59,
346-420.  This module is that graph with every arithmetic node, forward AND backward, executed by a CUDA kernel of
csrc/train.cu through the C ABI of include/b200sam_train.h:

  * each primitive below is a torch.autograd.Function whose forward and backward are ctypes calls; torch.autograd is
    only the tape (it remembers the order and hands the gradients back);
  * activations are token-major / channels-last fp32, heads and transposes are stride patterns of the strided GEMM
    (views here, never copies);
  * all prompts of all frames of the batch go through one decoder pass per side (the reference loops per image).

Nothing here runs without the CUDA library: there is no PyTorch implementation of any node to fall back to.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._manifest import GRID, OUT_CH

_I64x4 = C.c_int64 * 4
TOKENS = GRID * GRID


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def _f32(t: torch.Tensor, what: str) -> torch.Tensor:
    if t.dtype != torch.float32 or not t.is_cuda:
        raise TypeError(f"{what}: expected a CUDA float32 tensor, got {t.dtype} on {t.device}")
    return t


class _Ctx:
    """The library handle the primitives report errors through (any handle of the device)."""
    handle: Optional[C.c_void_p] = None

    @classmethod
    def h(cls) -> C.c_void_p:
        if cls.handle is None:
            raise RuntimeError("train primitives used before a B200Sam handle was bound (call B200Sam.train_ops())")
        return cls.handle


def bind_handle(handle: C.c_void_p) -> None:
    _Ctx.handle = handle


def _check(rc: int, what: str) -> None:
    _lib.check(rc, _Ctx.h(), what)


# ------------------------------------------------------------------------------------------------ raw launches
def _gemm(a: torch.Tensor, sa, b: torch.Tensor, sb, c: torch.Tensor, sc, nb1: int, nb2: int, m: int, n: int, k: int,
          alpha: float = 1.0, beta: float = 0.0, bias: Optional[torch.Tensor] = None, bias_mode: int = 0,
          relu: bool = False) -> None:
    ev = None
    if GEMM_PROFILE is not None:      # measurement aid (scripts/prof_train.py): device time per problem shape
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record()
    rc = _lib.load().b200sam_train_gemm(_Ctx.h(), _p(a), _I64x4(*sa), _p(b), _I64x4(*sb), _p(c), _I64x4(*sc), _p(bias),
                                        bias_mode, nb1, nb2, m, n, k, C.c_float(alpha), C.c_float(beta), int(relu),
                                        _stream())
    _check(rc, "train_gemm")
    if ev is not None:
        ev[1].record()
        GEMM_PROFILE.append(((nb1 * nb2, m, n, k, tuple(sa[2:]), tuple(sb[2:]), tuple(sc[2:])), ev))


# Arithmetic of the big Linear layers (rows >= TC_MIN_ROWS: the image-side projections over P * 4096 tokens and the two
# ConvTranspose2d).  "tf32": tcgen05 tensor cores, TF32 operands, fp32 accumulation - above the reference's own training
# precision (train_ds.py runs bf16 autocast) - for the forward, dX and dW.  "fp32": the exact SIMT kernel for
# everything (bit-for-bit reproducible forward; used by the tight-tolerance parity tests).  Token-side layers (a few
# rows) always run in fp32.
_PRECISION = "tf32"
TC_MIN_ROWS = 1024
GEMM_PROFILE: Optional[list] = None


def set_precision(mode: str) -> str:
    """Select "tf32" (default) or "fp32" for the big Linear layers; returns the previous mode."""
    global _PRECISION
    if mode not in ("tf32", "fp32"):
        raise ValueError("precision must be 'tf32' or 'fp32'")
    prev, _PRECISION = _PRECISION, mode
    return prev


def _gemm_tc(a: torch.Tensor, w: torch.Tensor, out: torch.Tensor, bias: Optional[torch.Tensor] = None,
             split_k: bool = False) -> None:
    """out (M, N) = a (M, K) @ w (N, K)^T (+ bias) on the tensor cores; all row-major 2-D with unit column stride."""
    M, K = a.shape
    N = w.shape[0]
    assert w.shape[1] == K and out.shape == (M, N) and a.stride(1) == 1 and w.stride(1) == 1 and out.stride(1) == 1
    rc = _lib.load().b200sam_train_gemm_tc(_Ctx.h(), _p(a), a.stride(0), _p(w), w.stride(0), _p(out), out.stride(0),
                                           _p(bias), M, N, K, int(split_k), _stream())
    _check(rc, "train_gemm_tc")


def _transpose(x: torch.Tensor) -> torch.Tensor:
    """(rows, cols) contiguous -> (cols, rows) contiguous."""
    rows, cols = x.shape
    assert x.is_contiguous()
    y = torch.empty((cols, rows), dtype=torch.float32, device=x.device)
    _check(_lib.load().b200sam_train_transpose(_Ctx.h(), _p(x), rows, cols, _p(y), _stream()), "train_transpose")
    return y


def _tc_ok(rows: int, K: int, N: int, *tensors: torch.Tensor) -> bool:
    return (_PRECISION == "tf32" and rows >= TC_MIN_ROWS and rows % 8 == 0 and K % 32 == 0 and N % 32 == 0
            and all(t.data_ptr() % 16 == 0 for t in tensors))


def _mm4(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, alpha: float = 1.0) -> None:
    """out (b1,b2,m,n) = alpha * a (b1,b2,m,k) @ b (b1,b2,k,n); all three are arbitrary-stride 4-D views."""
    b1, b2, m, k = a.shape
    n = b.shape[3]
    assert b.shape[:3] == (b1, b2, k) and out.shape == (b1, b2, m, n), (a.shape, b.shape, out.shape)
    _gemm(a, a.stride(), b, b.stride(), out, out.stride(), b1, b2, m, n, k, alpha)


def _empty_like_layout(t: torch.Tensor) -> torch.Tensor:
    """Uninitialised tensor with t's shape AND strides when t is a dense permutation (so that the backward of the
    view ops that produced t is again a pure view), else contiguous."""
    if t.is_contiguous() or t.numel() == 0:
        return torch.empty_like(t, memory_format=torch.contiguous_format)
    order = sorted(range(t.dim()), key=lambda d: (-t.stride(d), d))
    expect, dense = 1, True
    for d in reversed(order):
        if t.shape[d] != 1 and t.stride(d) != expect:
            dense = False
            break
        expect *= t.shape[d]
    if not dense:
        return torch.empty(t.shape, dtype=t.dtype, device=t.device)
    return torch.empty_strided(t.shape, t.stride(), dtype=t.dtype, device=t.device)


# ------------------------------------------------------------------------------------------------ primitives
class LinearFn(torch.autograd.Function):
    """y = x W^T + b (nn.Linear) or, with w_kn=True, y = x W + b for a weight stored (K, N) - ConvTranspose2d's
    (C_in, C_out*2*2) flat weight.  Optional fused ReLU (MLP blocks).  x: (..., K) with contiguous rows."""

    @staticmethod
    def forward(ctx, x, w, b, relu: bool, w_kn: bool):
        _f32(x, "linear x"), _f32(w, "linear w")
        K = x.shape[-1]
        N = w.shape[1] if w_kn else w.shape[0]
        assert (w.shape[0] if w_kn else w.shape[1]) == K and x.stride(-1) == 1 and w.is_contiguous()
        x2 = x.reshape(-1, K)
        assert x2.data_ptr() == x.data_ptr(), "linear input must be viewable as rows without a copy"
        rows = x2.shape[0]
        y = torch.empty((rows, N), dtype=torch.float32, device=x.device)
        tc = (not relu) and x2.stride(0) == K and _tc_ok(rows, K, N, x2, w)
        if tc:
            _gemm_tc(x2, _transpose(w) if w_kn else w, y, b)
        else:
            sb = (0, 0, N, 1) if w_kn else (0, 0, 1, K)
            _gemm(x2, (0, 0, x2.stride(0), 1), w, sb, y, (0, 0, N, 1), 1, 1, rows, N, K, bias=b,
                  bias_mode=1 if b is not None else 0, relu=relu)
        ctx.save_for_backward(x2, w, y if relu else None)
        ctx.cfg = (relu, w_kn, tuple(x.shape[:-1]), b is not None, tc)
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, w, y = ctx.saved_tensors
        relu, w_kn, lead, has_b, tc = ctx.cfg
        rows, K = x2.shape
        N = w.shape[1] if w_kn else w.shape[0]
        dy = dy.reshape(rows, N)
        if not dy.is_contiguous():
            dy = dy.contiguous()
        if tc and dy.data_ptr() % 16:
            dy = dy.clone()
        if relu:
            d2 = torch.empty_like(dy)
            _check(_lib.load().b200sam_train_act_bwd(_Ctx.h(), 0, _p(dy), _p(y), dy.numel(), _p(d2), _stream()),
                   "act_bwd")
            dy = d2
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((rows, K), dtype=torch.float32, device=dy.device)
            # dx = dy W (N,K)  /  dy W^T for the (K,N) storage
            if tc:      # the tensor-core kernel wants the second operand as (out features, contraction) rows
                _gemm_tc(dy, w if w_kn else _transpose(w), dx)
            else:
                sb = (0, 0, 1, N) if w_kn else (0, 0, K, 1)
                _gemm(dy, (0, 0, N, 1), w, sb, dx, (0, 0, K, 1), 1, 1, rows, K, N)
            dx = dx.view(*lead, K)
        if ctx.needs_input_grad[1]:
            dw = torch.empty_like(w)
            if tc:      # contraction over the rows: both operands transposed once, K split over the CTAs
                dyt, xt = _transpose(dy), _transpose(x2)
                if w_kn:
                    _gemm_tc(xt, dyt, dw, split_k=True)       # dW (K,N) = x^T dy
                else:
                    _gemm_tc(dyt, xt, dw, split_k=True)       # dW (N,K) = dy^T x
            elif w_kn:    # dW (K,N) = x^T dy
                _gemm(x2, (0, 0, 1, x2.stride(0)), dy, (0, 0, N, 1), dw, (0, 0, N, 1), 1, 1, K, N, rows)
            else:       # dW (N,K) = dy^T x
                _gemm(dy, (0, 0, 1, N), x2, (0, 0, x2.stride(0), 1), dw, (0, 0, K, 1), 1, 1, N, K, rows)
        if has_b and ctx.needs_input_grad[2]:
            db = torch.empty((N,), dtype=torch.float32, device=dy.device)
            _check(_lib.load().b200sam_train_colsum(_Ctx.h(), _p(dy), rows, N, _p(db), _stream()), "colsum")
        return dx, dw, db, None, None


def linear(x, w, b=None, relu: bool = False, w_kn: bool = False):
    return LinearFn.apply(x, w, b, relu, w_kn)


class MatMulFn(torch.autograd.Function):
    """out = alpha * a @ b on 4-D views (b1, b2, m, k) x (b1, b2, k, n).  `out_order` is the dimension order the
    result is laid out in memory (e.g. (0, 2, 1, 3) stores (b1, m, b2, n) contiguously = recombined heads)."""

    @staticmethod
    def forward(ctx, a, b, alpha: float, out_order: Tuple[int, int, int, int]):
        _f32(a, "matmul a"), _f32(b, "matmul b")
        b1, b2, m, _ = a.shape
        n = b.shape[3]
        shape = (b1, b2, m, n)
        phys = torch.empty([shape[d] for d in out_order], dtype=torch.float32, device=a.device)
        inv = [out_order.index(d) for d in range(4)]
        out = phys.permute(*inv)
        _mm4(a, b, out, alpha)
        ctx.save_for_backward(a, b)
        ctx.alpha = alpha
        return out

    @staticmethod
    def backward(ctx, dout):
        a, b = ctx.saved_tensors
        da = db = None
        if ctx.needs_input_grad[0]:
            da = _empty_like_layout(a)
            _mm4(dout, b.transpose(2, 3), da, ctx.alpha)
        if ctx.needs_input_grad[1]:
            db = _empty_like_layout(b)
            _mm4(a.transpose(2, 3), dout, db, ctx.alpha)
        return da, db, None, None


def matmul4(a, b, alpha: float = 1.0, out_order=(0, 1, 2, 3)):
    return MatMulFn.apply(a, b, alpha, tuple(out_order))


class LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, g, b, eps: float):
        _f32(x, "layernorm x")
        assert x.is_contiguous() and g.is_contiguous() and b.is_contiguous()
        Cn = x.shape[-1]
        rows = x.numel() // Cn
        y = torch.empty_like(x)
        mean = torch.empty((rows,), dtype=torch.float32, device=x.device)
        rstd = torch.empty_like(mean)
        _check(_lib.load().b200sam_train_layernorm_fwd(_Ctx.h(), _p(x), _p(g), _p(b), rows, Cn, C.c_float(eps), _p(y),
                                                       _p(mean), _p(rstd), _stream()), "layernorm_fwd")
        ctx.save_for_backward(x, g, mean, rstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, g, mean, rstd = ctx.saved_tensors
        Cn = x.shape[-1]
        rows = x.numel() // Cn
        dy = dy.contiguous()
        dx = torch.empty_like(x)
        dg = torch.empty_like(g)
        db = torch.empty_like(g)
        _check(_lib.load().b200sam_train_layernorm_bwd(_Ctx.h(), _p(dy), _p(x), _p(g), _p(mean), _p(rstd), rows, Cn,
                                                       _p(dx), _p(dg), _p(db), _stream()), "layernorm_bwd")
        return dx, dg, db, None


def layernorm(x, g, b, eps: float):
    return LayerNormFn.apply(x, g, b, eps)


class SoftmaxFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _f32(x, "softmax x")
        assert x.is_contiguous()
        L = x.shape[-1]
        y = torch.empty_like(x)
        _check(_lib.load().b200sam_train_softmax_fwd(_Ctx.h(), _p(x), x.numel() // L, L, _p(y), _stream()), "softmax_fwd")
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        L = y.shape[-1]
        dy = dy.contiguous()
        dx = torch.empty_like(y)
        _check(_lib.load().b200sam_train_softmax_bwd(_Ctx.h(), _p(dy), _p(y), y.numel() // L, L, _p(dx), _stream()),
               "softmax_bwd")
        return dx


def softmax(x):
    return SoftmaxFn.apply(x)


class ActFn(torch.autograd.Function):
    """kind 0 ReLU, 1 GELU (erf)."""

    @staticmethod
    def forward(ctx, x, kind: int):
        _f32(x, "act x")
        assert x.is_contiguous()
        y = torch.empty_like(x)
        _check(_lib.load().b200sam_train_act_fwd(_Ctx.h(), kind, _p(x), x.numel(), _p(y), _stream()), "act_fwd")
        ctx.save_for_backward(y if kind == 0 else x)
        ctx.kind = kind
        return y

    @staticmethod
    def backward(ctx, dy):
        (ref,) = ctx.saved_tensors
        dy = dy.contiguous()
        dx = torch.empty_like(ref)
        _check(_lib.load().b200sam_train_act_bwd(_Ctx.h(), ctx.kind, _p(dy), _p(ref), ref.numel(), _p(dx), _stream()),
               "act_bwd")
        return dx, None


def gelu(x):
    return ActFn.apply(x, 1)


def _add_rows(a: torch.Tensor, b: torch.Tensor, mod: int) -> torch.Tensor:
    Cn = a.shape[-1]
    rows = a.numel() // Cn
    out = torch.empty_like(a)
    _check(_lib.load().b200sam_train_add_bcast(_Ctx.h(), _p(a), _p(b), rows, Cn, mod, _p(out), _stream()), "add_bcast")
    return out


class AddFn(torch.autograd.Function):
    """a + b with b repeated along the rows with period `mod` rows (mod == rows(a): plain add).  The gradient of a
    broadcast b is not needed anywhere in the decoder (key_pe is a buffer) and is refused."""

    @staticmethod
    def forward(ctx, a, b, mod: int):
        _f32(a, "add a"), _f32(b, "add b")
        assert a.is_contiguous() and b.is_contiguous() and a.shape[-1] == b.shape[-1]
        rows = a.numel() // a.shape[-1]
        assert b.numel() == mod * a.shape[-1]
        ctx.plain = mod == rows
        ctx.bshape = b.shape
        return _add_rows(a, b, mod)

    @staticmethod
    def backward(ctx, g):
        db = None
        if ctx.needs_input_grad[1]:
            if not ctx.plain:
                raise RuntimeError("gradient of a broadcast addend is not implemented (not needed by the decoder)")
            db = g.reshape(ctx.bshape)
        return g, db, None


def add(a, b):
    return AddFn.apply(a, b, a.numel() // a.shape[-1])


def add_bcast(a, b, mod: int):
    return AddFn.apply(a, b, mod)


class ForkFn(torch.autograd.Function):
    """Explicit fan-out: returns n aliases of x; the backward sums the n gradients with the library's add kernel
    (otherwise autograd would sum them with its own)."""

    @staticmethod
    def forward(ctx, x, n: int):
        return tuple(x.view_as(x) for _ in range(n))

    @staticmethod
    def backward(ctx, *gs):
        acc = None
        for g in gs:
            if g is None:
                continue
            g = g.contiguous()
            acc = g if acc is None else _add_rows(acc, g, g.numel() // g.shape[-1])
        return acc, None


def fork(x, n: int):
    return ForkFn.apply(x, n)


class PixelShuffleFn(torch.autograd.Function):
    """(n, h, w, co*4) -> (n, 2h, 2w, co), column = co*4 + ky*2 + kx: the scatter half of ConvTranspose2d(k=2, s=2)."""

    @staticmethod
    def forward(ctx, x, n: int, hgt: int, wid: int, co: int):
        assert x.is_contiguous() and x.numel() == n * hgt * wid * co * 4
        y = torch.empty((n, 2 * hgt, 2 * wid, co), dtype=torch.float32, device=x.device)
        _check(_lib.load().b200sam_train_pixel_shuffle(_Ctx.h(), _p(x), n, hgt, wid, co, 0, _p(y), _stream()),
               "pixel_shuffle")
        ctx.cfg = (n, hgt, wid, co, tuple(x.shape))
        return y

    @staticmethod
    def backward(ctx, dy):
        n, hgt, wid, co, shape = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty(shape, dtype=torch.float32, device=dy.device)
        _check(_lib.load().b200sam_train_pixel_shuffle(_Ctx.h(), _p(dy), n, hgt, wid, co, 1, _p(dx), _stream()),
               "pixel_shuffle")
        return dx, None, None, None, None


class ResizeFn(torch.autograd.Function):
    """F.interpolate(x[..., :in_h, :in_w], (out_h, out_w), mode="bilinear", align_corners=False) on (n, H, W)."""

    @staticmethod
    def forward(ctx, x, in_h: int, in_w: int, out_h: int, out_w: int):
        _f32(x, "resize x")
        assert x.dim() == 3 and x.is_contiguous() and in_h <= x.shape[1] and in_w <= x.shape[2]
        n, H, W = x.shape
        y = torch.empty((n, out_h, out_w), dtype=torch.float32, device=x.device)
        _check(_lib.load().b200sam_train_resize_fwd(_Ctx.h(), _p(x), n, in_h, in_w, W, H * W, out_h, out_w, _p(y),
                                                    _stream()), "resize_fwd")
        ctx.cfg = (n, H, W, in_h, in_w, out_h, out_w)
        return y

    @staticmethod
    def backward(ctx, dy):
        n, H, W, in_h, in_w, out_h, out_w = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty((n, H, W), dtype=torch.float32, device=dy.device)
        _check(_lib.load().b200sam_train_resize_bwd(_Ctx.h(), _p(dy), n, in_h, in_w, W, H * W, out_h, out_w, _p(dx),
                                                    _stream()), "resize_bwd")
        return dx, None, None, None, None


class MaskLossFn(torch.autograd.Function):
    """(bce_mean[P], dice[P]) of LISA.py:16-59 per mask on weight[p] * logits[p]; logits, gt: (P, H, W) fp32."""

    SCALE, EPS = 1000.0, 1e-6    # dice_loss defaults, LISA.py:20-21

    @staticmethod
    def forward(ctx, logits, gt, weight):
        _f32(logits, "mask loss logits"), _f32(gt, "mask loss gt")
        assert logits.is_contiguous() and gt.is_contiguous() and logits.shape == gt.shape and logits.dim() == 3
        P = logits.shape[0]
        hw = logits.shape[1] * logits.shape[2]
        if weight is not None:
            weight = _f32(weight, "mask loss weight").contiguous()
            assert weight.numel() == P
        sums = torch.empty((P, 4), dtype=torch.float64, device=logits.device)
        _check(_lib.load().b200sam_train_mask_loss_fwd(_Ctx.h(), _p(logits), _p(gt), _p(weight), P, hw, _p(sums),
                                                       _stream()), "mask_loss_fwd")
        s, e = MaskLossFn.SCALE, MaskLossFn.EPS
        bce = (sums[:, 0] / hw).float()
        dice = (1.0 - (2.0 * sums[:, 2] / s + e) / (sums[:, 1] / s + sums[:, 3] / s + e)).float()
        ctx.save_for_backward(logits, gt, weight, sums)
        return bce, dice

    @staticmethod
    def backward(ctx, g_bce, g_dice):
        logits, gt, weight, sums = ctx.saved_tensors
        P = logits.shape[0]
        hw = logits.shape[1] * logits.shape[2]
        g_bce = g_bce.contiguous().float()
        g_dice = g_dice.contiguous().float()
        d = torch.empty_like(logits)
        _check(_lib.load().b200sam_train_mask_loss_bwd(_Ctx.h(), _p(logits), _p(gt), _p(weight), _p(sums), _p(g_bce),
                                                       _p(g_dice), P, hw, C.c_float(MaskLossFn.SCALE),
                                                       C.c_float(MaskLossFn.EPS), _p(d), _stream()), "mask_loss_bwd")
        return d, None, None


def src_tokens(emb_nchw: torch.Tensor, frame_of_prompt: Optional[torch.Tensor], dense_vec: torch.Tensor,
               n_prompts: int) -> torch.Tensor:
    """(P, 4096, 256) fp32: image_embeddings[frame_of_prompt[p]] + dense, token-major (no gradient: the encoder is
    frozen and the prompt encoder's no_mask_embed is not trained, LISA.py:81-83)."""
    dt = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16, torch.float16: _lib.F16}[emb_nchw.dtype]
    assert emb_nchw.is_contiguous() and tuple(emb_nchw.shape[1:]) == (OUT_CH, GRID, GRID)
    out = torch.empty((n_prompts, TOKENS, OUT_CH), dtype=torch.float32, device=emb_nchw.device)
    _check(_lib.load().b200sam_train_src_tokens(_Ctx.h(), _p(emb_nchw), dt, emb_nchw.shape[0], _p(frame_of_prompt),
                                                _p(_f32(dense_vec, "dense").contiguous()), n_prompts, _p(out),
                                                _stream()), "src_tokens")
    return out


# ------------------------------------------------------------------------------------------------ the decoder graph
HEADS = 8            # build_sam.py:88-96: TwoWayTransformer(depth=2, embedding_dim=256, mlp_dim=2048, num_heads=8)
DEPTH = 2
N_MASK_TOKENS = 4    # num_multimask_outputs + 1 (mask_decoder.py:44)

# Fan-out: the image-side tensor `keys` (P, 4096, 256) is consumed two or three times per block; those uses go through
# fork() so that the gradient sum is the library's add kernel.  The token-side tensors are (P, 6, 256) = a few KB;
# their fan-out sums, the cat / slice / expand that assemble and split the tokens, and dtype casts of non-fp32
# parameters are left to autograd's own bookkeeping.


class _Params:
    """fp32 views of one decoder's parameters by their state_dict names (bf16 / fp16 parameters are cast through
    autograd, so their .grad arrives in their own dtype)."""

    def __init__(self, dec: torch.nn.Module):
        self._p = dict(dec.named_parameters())

    def __call__(self, name: str) -> torch.Tensor:
        t = self._p[name]
        t = t if t.dtype == torch.float32 else t.float()
        return t if t.is_contiguous() else t.contiguous()

    def has(self, name: str) -> bool:
        return name in self._p


def _attention(p: _Params, pre: str, q, k, v):
    """transformer.Attention.forward (transformer.py:215-240).  q: (P, Tq, 256), k / v: (P, Tk, 256)."""
    P, Tq, _ = q.shape
    Tk = k.shape[1]
    qp = linear(q, p(pre + "q_proj.weight"), p(pre + "q_proj.bias"))
    kp = linear(k, p(pre + "k_proj.weight"), p(pre + "k_proj.bias"))
    vp = linear(v, p(pre + "v_proj.weight"), p(pre + "v_proj.bias"))
    inner = qp.shape[-1]
    d = inner // HEADS
    qh = qp.view(P, Tq, HEADS, d).permute(0, 2, 1, 3)          # (P, H, Tq, d)   _separate_heads
    kt = kp.view(P, Tk, HEADS, d).permute(0, 2, 3, 1)          # (P, H, d, Tk)   k.permute(0, 1, 3, 2)
    vh = vp.view(P, Tk, HEADS, d).permute(0, 2, 1, 3)          # (P, H, Tk, d)
    attn = softmax(matmul4(qh, kt, alpha=1.0 / math.sqrt(d)))  # (P, H, Tq, Tk)
    out = matmul4(attn, vh, out_order=(0, 2, 1, 3))            # stored (P, Tq, H, d) = _recombine_heads
    out = out.permute(0, 2, 1, 3).reshape(P, Tq, inner)        # a view of that storage
    return linear(out, p(pre + "out_proj.weight"), p(pre + "out_proj.bias"))


def _two_way_block(p: _Params, pre: str, queries, keys, query_pe, key_pe, skip_first_layer_pe: bool):
    """TwoWayAttentionBlock.forward (transformer.py:151-186)."""
    def ln(x, name):
        return layernorm(x, p(pre + name + ".weight"), p(pre + name + ".bias"), 1e-5)

    if skip_first_layer_pe:
        queries = _attention(p, pre + "self_attn.", queries, queries, queries)
    else:
        q = add(queries, query_pe)
        queries = add(queries, _attention(p, pre + "self_attn.", q, q, queries))
    queries = ln(queries, "norm1")

    # cross attention, tokens attending to the image
    k0, k1, keys = fork(keys, 3)
    q = add(queries, query_pe)
    k = add_bcast(k0, key_pe, TOKENS)
    queries = ln(add(queries, _attention(p, pre + "cross_attn_token_to_image.", q, k, k1)), "norm2")

    # MLP block (transformer.py MLPBlock: lin1 -> ReLU -> lin2)
    hid = linear(queries, p(pre + "mlp.lin1.weight"), p(pre + "mlp.lin1.bias"), relu=True)
    queries = ln(add(queries, linear(hid, p(pre + "mlp.lin2.weight"), p(pre + "mlp.lin2.bias"))), "norm3")

    # cross attention, image attending to the tokens
    k0, k1 = fork(keys, 2)
    q = add(queries, query_pe)
    k = add_bcast(k0, key_pe, TOKENS)
    keys = ln(add(k1, _attention(p, pre + "cross_attn_image_to_token.", k, q, queries)), "norm4")
    return queries, keys


def _mlp(p: _Params, pre: str, x, n_layers: int = 3):
    """mask_decoder.MLP (mask_decoder.py:183-207): ReLU between the layers, none after the last."""
    for i in range(n_layers):
        x = linear(x, p(f"{pre}layers.{i}.weight"), p(f"{pre}layers.{i}.bias"), relu=i < n_layers - 1)
    return x


def decoder_forward(dec: torch.nn.Module, emb_nchw: torch.Tensor, frame_of_prompt: Optional[torch.Tensor],
                    sparse: torch.Tensor, dense_vec: torch.Tensor, key_pe_tokens: torch.Tensor):
    """MaskDecoder.forward with multimask_output=False (mask_decoder.py:79-178) for P prompts over any number of
    frames.  emb_nchw: (F, 256, 64, 64) frozen image embeddings; frame_of_prompt: (P,) int32 or None when F == 1;
    sparse: (P, 1, 256) prompt embeddings (gradient flows back to the caller's text_hidden_fcs); dense_vec: (256,)
    no_mask_embed; key_pe_tokens: (4096, 256) dense PE in token order.
    Returns (masks (P, 1, 256, 256), iou_pred (P, 1), taxonomy (P, 4) or None)."""
    p = _Params(dec)
    P = sparse.shape[0]
    sparse = sparse if sparse.dtype == torch.float32 else sparse.float()
    # mask_decoder.py:131-138: tokens = [iou_token, mask_tokens, sparse]
    out_tok = torch.cat([p("iou_token.weight"), p("mask_tokens.weight")], 0)
    tokens = torch.cat([out_tok.unsqueeze(0).expand(P, -1, -1), sparse.reshape(P, -1, OUT_CH)], 1).contiguous()
    keys = src_tokens(emb_nchw, frame_of_prompt, dense_vec, P)          # (P, 4096, 256), no grad
    # TwoWayTransformer.forward (transformer.py:61-108); the first use of `keys` needs no gradient, every later one does
    queries = tokens
    for i in range(DEPTH):
        queries, keys = _two_way_block(p, f"transformer.layers.{i}.", queries, keys, tokens, key_pe_tokens, i == 0)
    k0, k1, keys = fork(keys, 3)
    q = add(queries, tokens)
    k = add_bcast(k0, key_pe_tokens, TOKENS)
    attn_out = _attention(p, "transformer.final_attn_token_to_image.", q, k, k1)
    hs = layernorm(add(queries, attn_out), p("transformer.norm_final_attn.weight"),
                   p("transformer.norm_final_attn.bias"), 1e-5)

    # output_upscaling (mask_decoder.py:52-61) on channels-last rows: ConvT(256->64) -> LayerNorm2d -> GELU ->
    # ConvT(64->32) -> GELU; ConvTranspose2d(k=2, s=2) = Linear with the (C_in, C_out*4) flat weight + pixel shuffle
    w0, b0 = p("output_upscaling.0.weight"), p("output_upscaling.0.bias")
    w3, b3 = p("output_upscaling.3.weight"), p("output_upscaling.3.bias")
    c1, c2 = w0.shape[1], w3.shape[1]
    y = linear(keys, w0.view(OUT_CH, c1 * 4), b0.repeat_interleave(4), w_kn=True)            # (P, 4096, 64*4)
    y = PixelShuffleFn.apply(y, P, GRID, GRID, c1)                                             # (P, 128, 128, 64)
    y = gelu(layernorm(y, p("output_upscaling.1.weight"), p("output_upscaling.1.bias"), 1e-6))
    y = linear(y, w3.view(c1, c2 * 4), b3.repeat_interleave(4), w_kn=True)                    # (P, 128, 128, 32*4)
    y = gelu(PixelShuffleFn.apply(y, P, 2 * GRID, 2 * GRID, c2))                               # (P, 256, 256, 32)
    side = 4 * GRID
    up = y.view(P, 1, side * side, c2)

    # hypernetworks (mask_decoder.py:153-163) and the mask product hyper_in @ upscaled
    hyper = torch.stack([_mlp(p, f"output_hypernetworks_mlps.{i}.", hs[:, 1 + i, :]) for i in range(N_MASK_TOKENS)], 1)
    masks = matmul4(hyper.view(P, 1, N_MASK_TOKENS, c2).contiguous(), up.permute(0, 1, 3, 2))  # (P, 1, 4, 65536)
    masks = masks.view(P, N_MASK_TOKENS, side, side)
    iou_pred = _mlp(p, "iou_prediction_head.", hs[:, 0, :])
    taxonomy = None
    if p.has("taxonomy_embed.layers.0.weight"):     # mask_decoder.py:117-118,172-178 (left decoder only)
        flat = hs[:, 1:1 + N_MASK_TOKENS, :].reshape(P, N_MASK_TOKENS * OUT_CH)
        taxonomy = softmax(_mlp(p, "taxonomy_embed.", flat.contiguous()))
    return masks[:, 0:1], iou_pred[:, 0:1], taxonomy          # mask_slice = slice(0, 1), mask_decoder.py:110-115


def postprocess_masks(masks: torch.Tensor, input_size: Sequence[int], original_size: Sequence[int],
                      img_size: int = 1024) -> torch.Tensor:
    """Sam.postprocess_masks (sam.py:155-189) under autograd: bilinear to img_size^2, crop to input_size, bilinear to
    original_size.  masks: (P, C, h, w) fp32 -> (P, C, H, W)."""
    P, Cn, hh, ww = masks.shape
    x = masks.reshape(P * Cn, hh, ww)
    x = x if x.is_contiguous() else x.contiguous()
    x = ResizeFn.apply(x, hh, ww, img_size, img_size)
    x = ResizeFn.apply(x, int(input_size[0]), int(input_size[1]), int(original_size[0]), int(original_size[1]))
    return x.view(P, Cn, int(original_size[0]), int(original_size[1]))


def mask_losses(pred: torch.Tensor, gt: torch.Tensor, weight: Optional[torch.Tensor] = None):
    """Per-mask (sigmoid_ce_loss mean over pixels, dice_loss term) of LISA.py:16-59 on weight[p] * pred[p].
    pred, gt: (P, H, W).  LISA.py:389-407 sums them per image and divides by the number of masks."""
    gt = gt if gt.dtype == torch.float32 else gt.float()
    return MaskLossFn.apply(pred.contiguous(), gt.contiguous(), weight)
