"""Helpers shared by the GPU parity tests (all calls go through the C ABI of libb200sam.so)."""
import ctypes as C
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG = "2handedafforder_b200"

# fp32 references computed with torch on the GPU must really be fp32 (SURVEY.md 8c hygiene note): cuDNN convolutions
# default to TF32, and matmul precision is pinned explicitly rather than relying on the default.
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
torch.set_float32_matmul_precision("highest")


def pkg():
    if os.path.exists(libmod().lib_path()):
        lib()                    # applies B200SAM_TEST_VARIANTS once
    return importlib.import_module(PKG)


_VARIANTS_APPLIED = False


def lib():
    """The loaded library.  B200SAM_TEST_VARIANTS="5=1,0=9" (kernel variant switches of include/b200sam_debug.h) runs
    the whole suite under non-default kernels, e.g. the folded LayerNorm."""
    global _VARIANTS_APPLIED
    L = importlib.import_module(PKG + "._lib").load()
    if not _VARIANTS_APPLIED:
        _VARIANTS_APPLIED = True
        for kv in filter(None, os.environ.get("B200SAM_TEST_VARIANTS", "").split(",")):
            k, v = kv.split("=")
            assert L.b200sam_debug_set_variant(int(k), int(v)) == 0
    return L


def libmod():
    return importlib.import_module(PKG + "._lib")


def P(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def rel_err(a: torch.Tensor, ref: torch.Tensor) -> float:
    a, ref = a.float(), ref.float()
    return float((a - ref).abs().max() / ref.abs().max().clamp_min(1e-30))


def op_gemm(A, W, mode, tf32=False, bias=None, addend=None, add_mod=0, ln_g=None, ln_b=None, eps=0.0,
            out=None, rowmap=0, simt=False):
    """C = A @ W^T through b200sam_op_gemm; allocates `out` if needed."""
    M, K = A.shape
    N = W.shape[0]
    if out is None:
        dt = torch.bfloat16 if mode in (0, 1) else torch.float32
        out = torch.empty((M, N), dtype=dt, device=A.device)
    rc = lib().b200sam_op_gemm(
        A.device.index or 0, int(tf32), mode, P(A), A.stride(0), P(W), W.stride(0), M, N, K, P(bias), P(addend),
        0 if addend is None else addend.stride(0), add_mod, P(ln_g), P(ln_b), float(eps), P(out), out.stride(0),
        rowmap, int(simt), stream())
    assert rc == 0, f"b200sam_op_gemm status {rc}"
    torch.cuda.synchronize()
    return out


def build_sam_from_synth(arch_name, **kw):
    """Facade with the seeded synthetic weights of oracle/synth.py loaded through load_state_dict."""
    from oracle import synth
    arch = synth.ARCHS[arch_name]
    sam = pkg()._build_sam(arch.dim, arch.depth, arch.heads, list(arch.global_idx), **kw)
    sd = synth.make_state_dict(arch, seed=0)
    missing, unexpected = sam.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    return sam.cuda(), arch, sd
