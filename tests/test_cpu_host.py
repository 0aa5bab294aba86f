"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol include/b200sam.h
declares, the facade keeps the reference's state_dict surface, and nothing silently falls back to the CPU."""
import ctypes
import importlib
import os
import re

import pytest
import torch

from oracle import synth
from tests import gpu_util as G

HEADERS = [os.path.join(G.ROOT, "include", n) for n in ("b200sam.h", "b200sam_debug.h", "b200sam_train.h")]


def _declared_functions():
    out = set()
    for path in HEADERS:
        src = open(path).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        out |= set(re.findall(r"\b(b200sam_[a-z0-9_]+)\s*\(", src))
    return sorted(out)


def test_library_exports_every_declared_symbol():
    libmod = G.libmod()
    if not os.path.exists(libmod.lib_path()):
        importlib.import_module(G.PKG + "._build").build_library()
    lib = ctypes.CDLL(libmod.lib_path())
    declared = _declared_functions()
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/*.h but not exported"
    bound = {n for n, _, _ in libmod.SYMBOLS}
    assert bound == set(declared), f"ctypes binding and header disagree: {bound ^ set(declared)}"
    libmod.load()


def test_create_without_gpu_fails_loudly():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = G.lib()
    h = ctypes.c_void_p()
    rc = lib.b200sam_create(ctypes.byref(h), 2, 0, 1, 3)
    assert rc != 0
    assert b"no CPU fallback" in lib.b200sam_last_error(h) or b"CUDA" in lib.b200sam_last_error(h)
    lib.b200sam_destroy(h)


@pytest.mark.parametrize("arch_name", ["tiny64", "vit_b", "vit_h"])
def test_facade_state_dict_matches_reference_manifest(arch_name):
    """Names/shapes equal the reference's state_dict (the synthetic dict is verified against the reference by
    load_state_dict(strict=True) in tests/golden/make_golden.py)."""
    arch = synth.ARCHS[arch_name]
    man = importlib.import_module(G.PKG + "._manifest").manifest(arch.dim, arch.depth, arch.heads, arch.global_idx)
    want = {k: tuple(v.shape) for k, v in synth.make_state_dict(synth.ARCHS["tiny64"], 0).items()} \
        if arch_name == "tiny64" else None
    names = [n for n, _, _ in man]
    assert len(names) == len(set(names))
    if arch_name == "vit_h":
        assert len(names) == 720
        assert sum(int(torch.Size(s).numel()) for _, s, b in man if not b) == 647_252_248   # Appendix B
    if want is not None:
        assert {n: s for n, s, _ in man} == want


def test_facade_load_roundtrip_and_param_filters():
    sam = G.pkg()._build_sam(256, 2, 4, [1])
    sd = synth.make_state_dict(synth.ARCHS["tiny64"], seed=0)
    res = sam.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    back = sam.state_dict()
    assert set(back) == set(sd)
    assert all(torch.equal(back[k], sd[k]) for k in sd)
    # train_ds.py:236-244 selects trainable params by substring
    dec = [n for n, _ in sam.named_parameters() if "mask_decoder" in n]
    assert any(n.startswith("mask_decoder_left.taxonomy_embed") for n in dec)
    assert not any(n.startswith("mask_decoder_right.taxonomy_embed") for n in dec)
    assert "prompt_encoder.pe_layer.positional_encoding_gaussian_matrix" in dict(sam.named_buffers())
    assert sam.image_encoder.img_size == 1024 and not sam.training
    sam = sam.bfloat16()
    assert sam.image_encoder.blocks._modules["0"].attn.qkv.weight.dtype == torch.bfloat16


def test_sam_pretrained_checkpoint_remap(tmp_path):
    """build_sam.py:125-136: `mask_decoder.*` of a raw SAM checkpoint feeds both sides."""
    sd = synth.make_state_dict(synth.ARCHS["tiny64"], seed=0)
    raw = {}
    for k, v in sd.items():
        if k.startswith("mask_decoder_left.") and "taxonomy" not in k:
            raw[k.replace("mask_decoder_left", "mask_decoder", 1)] = v
        elif not k.startswith("mask_decoder_"):
            raw[k] = v
    path = tmp_path / "sam.pth"
    torch.save(raw, path)
    sam = G.pkg()._build_sam(256, 2, 4, [1], checkpoint=str(path), is_sam_pretrained=True)
    got = sam.state_dict()
    k = "transformer.layers.1.cross_attn_image_to_token.k_proj.weight"
    assert torch.equal(got["mask_decoder_left." + k], sd["mask_decoder_left." + k])
    assert torch.equal(got["mask_decoder_right." + k], sd["mask_decoder_left." + k])


def test_no_cpu_fallback_in_facade():
    sam = G.pkg()._build_sam(256, 2, 4, [1])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sam.image_encoder(torch.zeros((1, 3, 1024, 1024)))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sam.postprocess_masks(torch.zeros((1, 1, 256, 256)), (576, 1024), (10, 10))


def test_product_package_never_imports_oracle():
    pkg_dir = os.path.join(G.ROOT, G.PKG)
    for root, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                assert "oracle" not in open(os.path.join(root, f)).read().replace("the oracle", ""), f


def test_training_module_has_no_cpu_path():
    """2handedafforder_b200/train.py composes CUDA primitives only: CPU tensors are refused by type, and nothing runs
    before a library handle is bound (no PyTorch implementation of any node to fall back to)."""
    T = importlib.import_module(G.PKG + ".train")
    x, w = torch.randn(4, 32), torch.randn(8, 32)
    with pytest.raises(TypeError, match="CUDA float32"):
        T.linear(x, w)
    with pytest.raises(TypeError, match="CUDA float32"):
        T.softmax(x)
    with pytest.raises(ValueError):
        T.set_precision("bf16")
    assert T.set_precision("fp32") in ("tf32", "fp32")
    assert T.set_precision("tf32") == "fp32"
    if not torch.cuda.is_available():
        saved = T._Ctx.handle
        T._Ctx.handle = None
        try:
            with pytest.raises(RuntimeError, match="before a B200Sam handle was bound"):
                T._check(0, "x")
        finally:
            T._Ctx.handle = saved
    # layout helper: a dense permutation keeps its strides, anything else becomes contiguous
    t = torch.empty(2, 3, 4, 5).permute(0, 2, 3, 1)
    assert T._empty_like_layout(t).stride() == t.stride()
    assert T._empty_like_layout(torch.empty(4, 6)[:, ::2]).is_contiguous()


def test_launch_summariser_classifies_the_pair_gemm_instantiations(tmp_path):
    """scripts/summarize_launches.py turns the ncu launch list into profiles/ncu_traffic.json (bench.py's
    roofline.traffic).  The pair GEMM's demangled names carry two template arguments since the epilogue got a
    compile-time leading dimension: proj follows an attention kernel, lin2 follows lin1, the remaining launch of the
    residual-stream mode is the patch embedding."""
    import json
    import subprocess
    import sys

    def rows(i, name, t_ns, rd, wr):
        base = [str(i), "1", "python", "host", name, "1", "7", "(384, 1, 1)", "(148, 1, 1)", "0", "10.0", "sec"]
        return [base + ["dram__bytes_read.sum", "byte", str(rd)], base + ["dram__bytes_write.sum", "byte", str(wr)],
                base + ["gpu__time_duration.sum", "ns", str(t_ns)]]

    pre = "void bsam::<unnamed>::"
    tail = "(CUtensorMap_st, CUtensorMap_st, int, int, int, bsam::GemmEpi)"
    seq = [("gemm_tc2_kernel<7, 1280>", 50000, 10, 20),                                                   # patch
           ("gemm_tc2_kernel<0, 0>", 200000, 100, 200),                                                     # qkv
           ("attn_window_tc_kernel_v1<80, 0>", 100000, 300, 60),
           ("gemm_tc2_kernel<7, 1280>", 110000, 250, 200),                                                  # proj
           ("gemm_tc2_kernel<1, 0>", 300000, 100, 290),                                                     # lin1
           ("gemm_tc2_kernel<7, 1280>", 290000, 600, 270)]                                                  # lin2
    src = tmp_path / "launches.csv"
    with open(src, "w") as fh:
        fh.write("==PROF== noise\n")
        fh.write('"ID","Process ID","Process Name","Host Name","Kernel Name","Context","Stream","Block Size","Grid Size",'
                 '"Device","CC","Section Name","Metric Name","Metric Unit","Metric Value"\n')
        for i, (n, t, rd, wr) in enumerate(seq):
            for r in rows(i, pre + n + tail, t, rd, wr):
                fh.write(",".join('"%s"' % c for c in r) + "\n")
    out = tmp_path / "out"
    out.mkdir()
    subprocess.run([sys.executable, os.path.join(G.ROOT, "scripts", "summarize_launches.py"), str(src), str(out), "t"],
                   check=True, capture_output=True)
    traffic = json.load(open(out / "ncu_traffic.json"))
    assert traffic["gemm_qkv"] == 300 and traffic["gemm_lin1"] == 390
    assert traffic["gemm_proj"] == 450 and traffic["gemm_lin2"] == 870 and traffic["gemm_patch"] == 30
    assert traffic["attn_window"] == 360
    assert os.path.exists(out / "t_launch_summary.csv")
