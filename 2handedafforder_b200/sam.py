"""Drop-in `visual_model` for 2HandedAfforder's LISA wrapper, backed by libb200sam (sm_100a CUDA).

Mirrors the reference `Sam` container as LISA drives it
(2Haff/model/segment_anything/modeling/sam.py:18-55,155-189; call sites 2Haff/model/LISA.py:157-168,
284-325, 487-532): submodules `image_encoder`, `prompt_encoder`, `mask_decoder_left`, `mask_decoder_right`
and the method `postprocess_masks`, with the reference's argument names, shapes and state_dict keys.
All arithmetic happens in the CUDA library; this file only validates arguments, owns the nn.Parameters
(so `from_pretrained`, `.bfloat16()`, `.cuda()`, `load_state_dict` keep working) and marshals pointers.

There is NO PyTorch / CPU fallback: calling any compute method without a CUDA device or without the built
library raises.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import nn

from . import _lib
from ._manifest import GRID, OUT_CH, manifest
from .callers import CallersMixin

IMG_SIZE = 1024   # build_sam.py:68

_DTYPES = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16, torch.float16: _lib.F16}


def _stream_ptr() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


class _Tree(nn.Module):
    """Parameter container; children/parameters are attached by dotted state_dict name."""

    def attach(self, dotted: str, value: torch.Tensor, is_buffer: bool) -> None:
        node = self
        parts = dotted.split(".")
        for p in parts[:-1]:
            if p not in node._modules:
                node.add_module(p, _Tree())
            node = node._modules[p]
        if is_buffer:
            node.register_buffer(parts[-1], value)
        else:
            node.register_parameter(parts[-1], nn.Parameter(value))


def _init_tensor(name: str, shape: Tuple[int, ...], g: torch.Generator) -> torch.Tensor:
    """Random init in the spirit of the reference modules' defaults (values are normally overwritten by a
    checkpoint).  pos_embed / rel_pos_* start at zero as in image_encoder.py:70-74,232-233."""
    leaf = name.rsplit(".", 1)[-1]
    if "pos_embed" in name or "rel_pos" in name:
        return torch.zeros(shape)
    if "gaussian_matrix" in name or "_token" in name or "embed" in name.split(".")[-2]:
        return torch.randn(shape, generator=g)
    if leaf == "bias":
        return torch.zeros(shape)
    if len(shape) == 1:   # norm weight
        return torch.ones(shape)
    fan_in = int(math.prod(shape[1:])) if len(shape) > 1 else shape[0]
    bound = 1.0 / math.sqrt(max(fan_in, 1))
    return (torch.rand(shape, generator=g) * 2 - 1) * bound


class _Component(_Tree):
    def _bind(self, owner: "B200Sam") -> None:
        object.__setattr__(self, "_owner", owner)   # not a registered submodule: no module cycle


class ImageEncoderB200(_Component):
    """ImageEncoderViT (image_encoder.py:17-125).  `forward` accepts any batch (the reference's
    get_visual_embs loops B=1, LISA.py:157-168)."""

    img_size = IMG_SIZE   # read by the reference's postprocess_masks (sam.py:180)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self._owner._encode(x, want_output=True)


class PromptEncoderB200(_Component):
    """PromptEncoder, text-prompt path (prompt_encoder.py:140-186)."""

    embed_dim = OUT_CH
    image_embedding_size = (GRID, GRID)
    input_image_size = (IMG_SIZE, IMG_SIZE)

    def get_dense_pe(self) -> torch.Tensor:
        """(1,256,64,64) positional encoding (prompt_encoder.py:67-76), fp32, computed once on the device."""
        return self._owner._dense_pe()

    def forward(self, points=None, boxes=None, masks=None, text_embeds=None):
        if points is not None or boxes is not None or masks is not None:
            raise NotImplementedError(
                "B200 SAM branch implements the text-prompt path only (points/boxes/masks are never passed by "
                "2Haff/model/LISA.py:288-293,498-503)")
        if text_embeds is None:
            raise ValueError("text_embeds is required")
        if text_embeds.dim() != 3 or text_embeds.shape[1] != 1 or text_embeds.shape[2] != OUT_CH:
            raise ValueError(f"text_embeds must be (N,1,{OUT_CH}), got {tuple(text_embeds.shape)}")
        n = text_embeds.shape[0]
        # cat with an fp32 empty tensor promotes to fp32 (prompt_encoder.py:165-177, Appendix A9)
        sparse = text_embeds.to(torch.float32)
        w = self.no_mask_embed.weight
        dense = w.reshape(1, -1, 1, 1).expand(n, -1, GRID, GRID)   # stride-0 view (prompt_encoder.py:182-184)
        return sparse, dense


class MaskDecoderB200(_Component):
    """MaskDecoder (mask_decoder.py:16-178) of one side."""

    def __init__(self, side: str):
        super().__init__()
        self.side = side

    def forward(self, image_embeddings, image_pe=None, sparse_prompt_embeddings=None,
                dense_prompt_embeddings=None, multimask_output: bool = False):
        if multimask_output:
            raise NotImplementedError("multimask_output=True is never used by LISA (LISA.py:280,489)")
        return self._owner._decode_one_side(self.side, image_embeddings, sparse_prompt_embeddings,
                                            dense_prompt_embeddings)


class B200Sam(CallersMixin, nn.Module):
    """The `visual_model` drop-in.  See module docstring."""

    mask_threshold: float = 0.0    # sam.py:19
    image_format: str = "RGB"

    def __init__(self, encoder_embed_dim: int, encoder_depth: int, encoder_num_heads: int,
                 encoder_global_attn_indexes: Sequence[int], max_batch: int = 8, max_prompts: int = 32,
                 seed: int = 0):
        super().__init__()
        self.arch = (int(encoder_embed_dim), int(encoder_depth), int(encoder_num_heads),
                     tuple(int(i) for i in encoder_global_attn_indexes))
        self.image_encoder = ImageEncoderB200()
        self.prompt_encoder = PromptEncoderB200()
        self.mask_decoder_left = MaskDecoderB200("left")
        self.mask_decoder_right = MaskDecoderB200("right")
        for c in (self.image_encoder, self.prompt_encoder, self.mask_decoder_left, self.mask_decoder_right):
            c._bind(self)
        g = torch.Generator().manual_seed(seed)
        self._names: List[str] = []
        for name, shape, is_buf in manifest(*self.arch):
            top, rest = name.split(".", 1)
            getattr(self, top).attach(rest, _init_tensor(name, shape, g), is_buf)
            self._names.append(name)
        # non-persistent buffers of the reference (sam.py:48-51); callers normalise on the host
        self.register_buffer("pixel_mean", torch.tensor([123.675, 116.28, 103.53]).view(-1, 1, 1), False)
        self.register_buffer("pixel_std", torch.tensor([58.395, 57.12, 57.375]).view(-1, 1, 1), False)
        self._handle: Optional[C.c_void_p] = None
        self._handle_device: Optional[int] = None
        self._cap = (int(max_batch), int(max_prompts))
        self._sig = None
        self._frozen = False
        self._pe_cache: Optional[torch.Tensor] = None
        self._pe_tok_cache: Optional[Tuple[torch.Tensor, torch.Tensor]] = None
        self._pe_key: Optional[Tuple[int, int, str]] = None
        self.eval()

    # ------------------------------------------------------------------ handle / weights
    @property
    def device(self) -> torch.device:
        return self.pixel_mean.device

    def _tensors(self) -> List[Tuple[str, torch.Tensor]]:
        sd = dict(self.named_parameters())
        sd.update(dict(self.named_buffers()))
        return [(n, sd[n]) for n in self._names]

    def _destroy(self) -> None:
        if self._handle is not None:
            _lib.load().b200sam_destroy(self._handle)
            self._handle = None
            self._sig = None
            self._pe_cache = None

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    def _ensure(self, device: torch.device, batch: int = 0, prompts: int = 0, sync: bool = True) -> C.c_void_p:
        if device.type != "cuda":
            raise RuntimeError("B200Sam needs a CUDA (sm_100a) device: there is no CPU fallback")
        lib = _lib.load()
        idx = device.index if device.index is not None else torch.cuda.current_device()
        cap_b, cap_p = self._cap
        if batch > cap_b or prompts > cap_p:   # grow the pre-allocated workspace (outside the hot path)
            self._cap = (max(cap_b, batch), max(cap_p, prompts))
            self._destroy()
        if self._handle is not None and self._handle_device != idx:
            self._destroy()
        if self._handle is None:
            h = C.c_void_p()
            dim, depth, heads, gidx = self.arch
            arr = (C.c_int * len(gidx))(*gidx)
            rc = lib.b200sam_create_custom(C.byref(h), dim, depth, heads, arr, len(gidx), idx, self._cap[0],
                                           self._cap[1])
            if rc != 0:
                msg = _lib.last_error(h)
                lib.b200sam_destroy(h)
                raise RuntimeError(f"b200sam_create failed (status {rc}): {msg}")
            self._handle, self._handle_device = h, idx
            self._sig = None
        if sync and not (self._frozen and self._sig is not None):
            self._sync_weights()
        return self._handle

    def _ensure_handle_only(self, device: torch.device) -> C.c_void_p:
        """Handle for the entry points that read no SAM weights (preprocess / seg_project / mask_metrics)."""
        return self._ensure(device, sync=False)

    def _sync_weights(self) -> None:
        """Upload the tensors whose (storage, version, dtype) changed since the last call and re-finalise.  Only the
        changed tensors travel, and the library rebuilds the derived buffers of the touched group (encoder /
        decoders) in place of the old ones, so validate() between training epochs of the mask decoders
        (train_ds.py:236-244,625-758) costs a few MB of copies and no extra memory."""
        tensors = self._tensors()
        sig = tuple((t.data_ptr(), t._version, t.dtype) for _, t in tensors)
        if sig == self._sig:
            return
        old = self._sig if (self._sig is not None and len(self._sig) == len(sig)) else (None,) * len(sig)
        lib, h = _lib.load(), self._handle
        # load_weight / finalize run on the legacy stream: everything queued on the caller's stream (producers of
        # the parameter values, kernels still reading the previous weights) must be complete first
        torch.cuda.current_stream(tensors[0][1].device).synchronize()
        for (name, t), s_new, s_old in zip(tensors, sig, old):
            if s_new == s_old:
                continue
            src = t.detach()
            if src.dtype not in _DTYPES:
                src = src.float()
            src = src.contiguous()
            shape = (C.c_int64 * src.dim())(*src.shape)
            _lib.check(lib.b200sam_load_weight(h, name.encode(), _ptr(src), _DTYPES[src.dtype], shape, src.dim()),
                       h, f"load_weight({name})")
        _lib.check(lib.b200sam_finalize_weights(h), h, "finalize_weights")
        self._sig = sig

    def freeze_weights(self, frozen: bool = True) -> None:
        """Skip the per-call parameter-change check (inference servers / benchmarks)."""
        self._frozen = frozen

    def launch_count(self) -> int:
        return 0 if self._handle is None else int(_lib.load().b200sam_launch_count(self._handle))

    # ------------------------------------------------------------------ encode
    def _encode(self, x: torch.Tensor, want_output: bool) -> Optional[torch.Tensor]:
        if x.dim() != 4 or tuple(x.shape[1:]) != (3, IMG_SIZE, IMG_SIZE):
            raise ValueError(f"frames must be (B,3,{IMG_SIZE},{IMG_SIZE}), got {tuple(x.shape)}")
        if x.dtype not in _DTYPES:
            raise TypeError(f"unsupported frame dtype {x.dtype}")
        self._no_autograd("image_encoder", x)      # frozen in the reference (LISA.py:81-83): no backward pass
        h = self._ensure(x.device, batch=x.shape[0])
        x = x.contiguous()
        out = torch.empty((x.shape[0], OUT_CH, GRID, GRID), dtype=x.dtype, device=x.device) if want_output else None
        with torch.cuda.device(x.device):
            rc = _lib.load().b200sam_encode(h, _ptr(x), _DTYPES[x.dtype], x.shape[0], _ptr(out), _DTYPES[x.dtype],
                                            _stream_ptr())
        _lib.check(rc, h, "encode")
        return out

    @torch.no_grad()
    def get_visual_embs(self, pixel_values: torch.Tensor) -> torch.Tensor:
        """Batched equivalent of LISAForCausalLM.get_visual_embs (LISA.py:157-168)."""
        return self._encode(pixel_values, want_output=True)

    # ------------------------------------------------------------------ prompt encoder / decoder
    def _dense_pe(self) -> torch.Tensor:
        dev = self.device
        # the PE depends on one never-trained buffer only: as long as that buffer is unchanged the cached tensor is
        # returned without the weight-change check of _ensure (LISA's training loop asks for it once per image while
        # the decoder parameters change every step)
        g = self.prompt_encoder.pe_layer.positional_encoding_gaussian_matrix
        key = (g.data_ptr(), g._version, str(dev))
        if self._pe_cache is not None and self._pe_key == key:
            return self._pe_cache
        h = self._ensure(dev)
        out = torch.empty((1, OUT_CH, GRID, GRID), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.load().b200sam_dense_pe(h, _ptr(out), _stream_ptr()), h, "dense_pe")
        self._pe_cache, self._pe_key = out, key
        return self._pe_cache

    def _check_dense(self, dense: Optional[torch.Tensor]) -> None:
        if dense is None or dense.numel() == 0:
            return
        w = self.prompt_encoder.no_mask_embed.weight
        if dense.data_ptr() != w.data_ptr() or dense.stride(-1) != 0 or dense.stride(-2) != 0:
            raise NotImplementedError(
                "dense_prompt_embeddings must be the no-mask embedding returned by prompt_encoder(...) "
                "(mask prompts are not on the LISA path)")

    def _decode_raw(self, emb: Optional[torch.Tensor], n_frames: int, seg: torch.Tensor, offsets: Sequence[int],
                    left: bool, right: bool, device: torch.device,
                    heads_out: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        """heads_out: optional (3, cap >= P, 4) f32 buffer; taxonomy / iou_left / iou_right are then written straight
        into heads_out[0 / 1 / 2, :P] (e.g. the send slot of clip.segment_clip) instead of fresh tensors."""
        P = int(offsets[-1])
        f32 = dict(dtype=torch.float32, device=device)
        out: Dict[str, torch.Tensor] = {}
        self._no_autograd("the fused inference decode (segment / segment_bits / eval-mode decoders)", emb, seg)
        if heads_out is not None:
            if (heads_out.dim() != 3 or heads_out.shape[0] != 3 or heads_out.shape[1] < P or heads_out.shape[2] != 4
                    or heads_out.dtype != torch.float32 or not heads_out.is_contiguous() or heads_out.device != device):
                raise ValueError("heads_out must be a contiguous (3, >=P, 4) float32 tensor on the frames' device")
        if left:
            out["lowres_left"] = torch.empty((P, 256, 256), **f32)
            out["iou_left"] = heads_out[1, :P] if heads_out is not None else torch.empty((P, 4), **f32)
            out["taxonomy"] = heads_out[0, :P] if heads_out is not None else torch.empty((P, 4), **f32)
        if right:
            out["lowres_right"] = torch.empty((P, 256, 256), **f32)
            out["iou_right"] = heads_out[2, :P] if heads_out is not None else torch.empty((P, 4), **f32)
        if P == 0:
            return out
        h = self._ensure(device, batch=n_frames, prompts=P)
        seg = seg.to(device=device, dtype=torch.float32).contiguous()
        offs = (C.c_int * len(offsets))(*[int(o) for o in offsets])
        emb_dt = _lib.F32
        if emb is not None:
            if emb.dtype not in _DTYPES:
                raise TypeError(f"unsupported embedding dtype {emb.dtype}")
            emb = emb.contiguous()
            emb_dt = _DTYPES[emb.dtype]
        with torch.cuda.device(device):
            rc = _lib.load().b200sam_decode(
                h, _ptr(emb), emb_dt, n_frames, _ptr(seg), offs,
                _ptr(out.get("lowres_left")), _ptr(out.get("lowres_right")), _ptr(out.get("iou_left")),
                _ptr(out.get("iou_right")), _ptr(out.get("taxonomy")), _stream_ptr())
        _lib.check(rc, h, "decode")
        return out

    def _wants_grad(self, *tensors: Optional[torch.Tensor], module: Optional[nn.Module] = None) -> bool:
        """True when the call is part of a training forward: autograd is on and an input requires grad, or the
        decoder is in train() mode with requires_grad parameters (LISA.py:84-90 does exactly that).  Plain inference -
        eval() mode, inputs without grad - stays on the fused inference kernels even outside no_grad()."""
        if not torch.is_grad_enabled():
            return False
        if any(t is not None and t.requires_grad for t in tensors):
            return True
        return module is not None and module.training and any(p.requires_grad for p in module.parameters())

    def _no_autograd(self, what: str, *tensors: Optional[torch.Tensor], module: Optional[nn.Module] = None) -> None:
        """Entry points without a backward pass (the frozen encoder, the fused segment() fast path) refuse inputs that
        require grad instead of returning detached tensors into a graph that expects gradients."""
        if self._wants_grad(*tensors, module=module):
            raise NotImplementedError(
                f"{what}: called with autograd enabled on inputs/parameters that require grad, but this entry point "
                "has no backward pass.  The trainable part of the SAM branch (both mask decoders, postprocess_masks, "
                "the mask losses) has one: use mask_decoder_left / mask_decoder_right / postprocess_masks or "
                "train_decode(); wrap everything else in torch.no_grad() as LISA.get_visual_embs does.")

    # ------------------------------------------------------------------ training forward / backward
    def train_ops(self, device: Optional[torch.device] = None):
        """The training module (2handedafforder_b200/train.py) bound to this model's library handle."""
        from . import train as T
        T.bind_handle(self._ensure_handle_only(device if device is not None else self.device))
        return T

    def _key_pe_tokens(self) -> torch.Tensor:
        """Dense positional encoding in token order (4096, 256) fp32 (transformer.py:84: image_pe.flatten(2).permute)."""
        pe = self._dense_pe()
        if self._pe_tok_cache is None or self._pe_tok_cache[0] is not pe:
            with torch.no_grad():
                self._pe_tok_cache = (pe, pe.flatten(2).permute(0, 2, 1)[0].contiguous())
        return self._pe_tok_cache[1]

    def _decode_train(self, side: str, image_embeddings: torch.Tensor, frame_of_prompt: Optional[torch.Tensor],
                      sparse: torch.Tensor):
        if image_embeddings.requires_grad:
            raise NotImplementedError("image_embeddings require grad, but the image encoder is frozen in the reference "
                                      "(LISA.py:81-83) and has no backward pass here")
        dec = self.mask_decoder_left if side == "left" else self.mask_decoder_right
        T = self.train_ops(image_embeddings.device)
        key_pe = self._key_pe_tokens()
        dense_vec = self.prompt_encoder.no_mask_embed.weight.detach().reshape(-1).float()
        with torch.cuda.device(image_embeddings.device):
            return T.decoder_forward(dec, image_embeddings.contiguous(), frame_of_prompt, sparse, dense_vec, key_pe)

    def train_decode(self, image_embeddings: torch.Tensor, seg_embeds: Sequence[torch.Tensor]) -> Dict[str, object]:
        """Both decoders for every prompt of every frame in ONE pass per side, under autograd (the reference loops
        per image, LISA.py:284-325).  image_embeddings: (F, 256, 64, 64) from get_visual_embs (no grad); seg_embeds:
        F tensors (N_i, 256), e.g. text_hidden_fcs outputs that require grad.
        Returns low_res_left / low_res_right (P, 1, 256, 256), iou_left / iou_right (P, 1), taxonomy (P, 4), offsets."""
        if image_embeddings.dim() != 4 or tuple(image_embeddings.shape[1:]) != (OUT_CH, GRID, GRID):
            raise ValueError(f"image_embeddings must be (F,{OUT_CH},{GRID},{GRID}), got {tuple(image_embeddings.shape)}")
        if len(seg_embeds) != image_embeddings.shape[0]:
            raise ValueError("one seg_embeds tensor per frame is required")
        dev = image_embeddings.device
        counts = [int(s.shape[0]) for s in seg_embeds]
        offsets = [0]
        for c in counts:
            offsets.append(offsets[-1] + c)
        P = offsets[-1]
        sparse = torch.cat([s.reshape(-1, OUT_CH) for s in seg_embeds], 0).reshape(P, 1, OUT_CH)
        fop = torch.tensor([i for i, c in enumerate(counts) for _ in range(c)], dtype=torch.int32, device=dev)
        ml, il, tax = self._decode_train("left", image_embeddings, fop, sparse)
        mr, ir, _ = self._decode_train("right", image_embeddings, fop, sparse)
        return {"low_res_left": ml, "low_res_right": mr, "iou_left": il, "iou_right": ir, "taxonomy": tax,
                "offsets": offsets}

    def _decode_one_side(self, side: str, image_embeddings: torch.Tensor, sparse: torch.Tensor,
                         dense: Optional[torch.Tensor]):
        dec = self.mask_decoder_left if side == "left" else self.mask_decoder_right
        train = self._wants_grad(sparse, module=dec)
        if image_embeddings.dim() != 4 or tuple(image_embeddings.shape) != (1, OUT_CH, GRID, GRID):
            raise ValueError(f"image_embeddings must be (1,{OUT_CH},{GRID},{GRID}) per call (LISA.py:306,505), got "
                             f"{tuple(image_embeddings.shape)}")
        if sparse is None or sparse.dim() != 3 or sparse.shape[1] != 1 or sparse.shape[2] != OUT_CH:
            raise ValueError("sparse_prompt_embeddings must be (N,1,256): exactly one [SEG] embedding per prompt")
        self._check_dense(dense)
        n = sparse.shape[0]
        if train:      # LISA.model_forward with train_mask_decoder (LISA.py:84-90,296-321): fp32 kernels with a backward
            masks, iou, tax = self._decode_train(side, image_embeddings, None, sparse)
            return (masks, iou, tax) if side == "left" else (masks, iou)
        res = self._decode_raw(image_embeddings, 1, sparse.reshape(n, OUT_CH), [0, n], side == "left", side == "right",
                               image_embeddings.device)
        masks = res[f"lowres_{side}"].reshape(n, 1, 256, 256)
        iou = res[f"iou_{side}"][:, 0:1]                       # multimask_output=False slice (mask_decoder.py:110-115)
        if side == "left":
            return masks, iou, res["taxonomy"]
        return masks, iou

    # ------------------------------------------------------------------ postprocess
    def _postprocess_raw(self, lowres: torch.Tensor, input_size, original_size, want_logits: bool,
                         want_mask: bool, threshold: float):
        n = lowres.shape[0]
        h_in, w_in = int(input_size[0]), int(input_size[1])
        H, W = int(original_size[0]), int(original_size[1])
        dev = lowres.device
        logits = torch.empty((n, H, W), dtype=torch.float32, device=dev) if want_logits else None
        mask = torch.empty((n, H, W), dtype=torch.uint8, device=dev) if want_mask else None
        if n == 0:
            return logits, mask
        h = self._ensure(dev)
        with torch.cuda.device(dev):
            rc = _lib.load().b200sam_postprocess(h, _ptr(lowres), n, h_in, w_in, H, W, _ptr(logits), _ptr(mask),
                                                 float(threshold), _stream_ptr())
        _lib.check(rc, h, "postprocess")
        return logits, mask

    def postprocess_masks(self, masks: torch.Tensor, input_size: Tuple[int, ...],
                          original_size: Tuple[int, ...]) -> torch.Tensor:
        """Sam.postprocess_masks (sam.py:155-189): returns fp32 (N,C,H,W)."""
        if masks.dim() != 4 or tuple(masks.shape[-2:]) != (256, 256):
            raise ValueError(f"masks must be (N,C,256,256), got {tuple(masks.shape)}")
        if self._wants_grad(masks):     # LISA.py:304-321 in the training forward
            T = self.train_ops(masks.device)
            with torch.cuda.device(masks.device):
                return T.postprocess_masks(masks.float(), input_size, original_size, IMG_SIZE)
        n, c = masks.shape[:2]
        low = masks.to(torch.float32).contiguous().reshape(n * c, 256, 256)
        logits, _ = self._postprocess_raw(low, input_size, original_size, True, False, 0.0)
        return logits.reshape(n, c, int(original_size[0]), int(original_size[1]))

    # ------------------------------------------------------------------ 1-bit wire format
    @staticmethod
    def bits_row_bytes(width: int) -> int:
        return (int(width) + 7) // 8

    @staticmethod
    def unpack_bits(bits: torch.Tensor, width: int) -> torch.Tensor:
        """(..., H, ceil(W/8)) packed u8 -> (..., H, W) u8 {0,1}; numpy.packbits(axis=-1) bit order.  A host/device
        convenience for consumers of the wire format (tests, file writers); not used on the hot path."""
        shifts = torch.arange(7, -1, -1, device=bits.device, dtype=torch.uint8)
        full = (bits.unsqueeze(-1) >> shifts) & 1
        return full.reshape(*bits.shape[:-1], bits.shape[-1] * 8)[..., :width].contiguous()

    def checksum64(self, data: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Per-item 64-bit checksum (b200sam_checksum64) of a contiguous (n_items, ...) u8 tensor -> (n_items,) int64
        (the bit pattern of the unsigned sum)."""
        if data.dtype != torch.uint8 or not data.is_contiguous() or data.dim() < 1:
            raise ValueError("checksum64 needs a contiguous uint8 tensor (n_items, ...)")
        n = data.shape[0]
        nbytes = data[0].numel() if n else 0
        if out is None:
            out = torch.empty((n,), dtype=torch.int64, device=data.device)
        elif out.dtype != torch.int64 or out.numel() < n or not out.is_contiguous():
            raise ValueError("out must be a contiguous int64 tensor with one element per item")
        if n == 0:
            return out
        h = self._ensure_handle_only(data.device)
        with torch.cuda.device(data.device):
            rc = _lib.load().b200sam_checksum64(h, _ptr(data), n, nbytes, nbytes, _ptr(out), _stream_ptr())
        _lib.check(rc, h, "checksum64")
        return out

    @torch.no_grad()
    def segment_bits(self, frames: torch.Tensor, seg_embeds: Sequence[torch.Tensor], resize, original,
                     threshold: float = 0.0, bits_out: Optional[torch.Tensor] = None,
                     heads_out: Optional[torch.Tensor] = None, gt: Optional[torch.Tensor] = None,
                     gt_index: Optional[torch.Tensor] = None) -> Dict[str, object]:
        """`segment` for a batch that shares one (resize, original) size - a video clip - with the masks in the 1-bit
        wire format and no staging copies: both decoders' masks are resampled and thresholded in ONE pass
        (b200sam_postprocess_pair) straight into `bits_out` (2, cap >= P, H, ceil(W/8)) u8 [0 = left, 1 = right] and
        the heads into `heads_out` (3, cap, 4) f32 [taxonomy, iou_left, iou_right] - clip.segment_clip passes views of
        its NCCL send slot.  With `gt` ((G,H,W) u8) the evaluation of train_ds.py:662-709,761-796 is fused behind the
        threshold: `counts` (P,3) int64 {intersection, union, area} and `pred_bits` (P,H,ceil(W/8)) come back too, and
        no u8 plane is ever written.  Returns bits_left / bits_right (P,H,Wb) views, taxonomy, iou_left, iou_right,
        offsets."""
        B = frames.shape[0]
        if len(seg_embeds) != B:
            raise ValueError("seg_embeds must have one entry per frame")  # LISA.py:193
        offsets = [0]
        for e in seg_embeds:
            if e.dim() != 2 or e.shape[1] != OUT_CH:
                raise ValueError(f"seg_embeds entries must be (N,{OUT_CH})")
            offsets.append(offsets[-1] + e.shape[0])
        P = offsets[-1]
        dev = frames.device
        H, W = int(original[0]), int(original[1])
        Wb = self.bits_row_bytes(W)
        if bits_out is None:
            bits_out = torch.empty((2, P, H, Wb), dtype=torch.uint8, device=dev)
        elif (bits_out.dim() != 4 or bits_out.shape[0] != 2 or bits_out.shape[1] < P
              or tuple(bits_out.shape[2:]) != (H, Wb) or bits_out.dtype != torch.uint8
              or not bits_out.is_contiguous() or bits_out.device != dev):
            raise ValueError(f"bits_out must be a contiguous (2, >={P}, {H}, {Wb}) uint8 tensor on {dev}")
        self._ensure(dev, batch=B, prompts=P)
        self._encode(frames, want_output=False)
        seg = torch.cat([e.to(device=dev, dtype=torch.float32) for e in seg_embeds], 0) if B else \
            torch.empty((0, OUT_CH), device=dev)
        res = self._decode_raw(None, B, seg, offsets, True, True, dev, heads_out=heads_out)
        out: Dict[str, object] = {"taxonomy": res["taxonomy"], "iou_left": res["iou_left"],
                                  "iou_right": res["iou_right"], "offsets": offsets,
                                  "lowres_left": res["lowres_left"], "lowres_right": res["lowres_right"],
                                  "bits_left": bits_out[0, :P], "bits_right": bits_out[1, :P]}
        counts = pred = None
        if gt is not None:
            if gt.dtype != torch.uint8 or gt.dim() != 3 or tuple(gt.shape[1:]) != (H, W) or not gt.is_contiguous():
                raise ValueError(f"gt must be a contiguous (G,{H},{W}) uint8 tensor")
            if gt_index is not None:
                gt_index = gt_index.to(device=dev, dtype=torch.int32).contiguous()
            elif gt.shape[0] < P:
                raise ValueError("gt has fewer masks than prompts and no gt_index was given")
            counts = torch.empty((P, 3), dtype=torch.int64, device=dev)
            pred = torch.empty((P, H, Wb), dtype=torch.uint8, device=dev)
            out["counts"], out["pred_bits"] = counts, pred
        if P:
            h = self._handle
            with torch.cuda.device(dev):
                rc = _lib.load().b200sam_postprocess_pair(
                    h, _ptr(res["lowres_left"]), _ptr(res["lowres_right"]), P, int(resize[0]), int(resize[1]), H, W,
                    _ptr(out["bits_left"]), _ptr(out["bits_right"]), float(threshold), _ptr(res["taxonomy"]),
                    _ptr(gt), _ptr(gt_index), _ptr(counts), _ptr(pred), _stream_ptr())
            _lib.check(rc, h, "postprocess_pair")
        return out

    # ------------------------------------------------------------------ fused fast path
    @torch.no_grad()
    def segment(self, frames: torch.Tensor, seg_embeds: Sequence[torch.Tensor], resize_list, original_size_list,
                threshold: float = 0.0, return_logits: bool = False) -> Dict[str, object]:
        """encode + both decoders + postprocess + threshold for a batch of frames, without per-image Python
        loops over kernels (the sequence LISA.evaluate runs per image, LISA.py:487-532).

        frames (B,3,1024,1024); seg_embeds: B tensors (N_i,256).  Returns a dict with
        `masks_left` / `masks_right`: list of B uint8 tensors (N_i,H_i,W_i) = logit > threshold,
        `logits_left` / `logits_right` (only if return_logits), `taxonomy` (P,4), `iou_left`, `iou_right` (P,4),
        `offsets` (B+1); `masks_left_packed` / `masks_right_packed` (P,H,W) when every frame shares one size.
        """
        B = frames.shape[0]
        if len(seg_embeds) != B or len(resize_list) != B or len(original_size_list) != B:
            raise ValueError("seg_embeds / resize_list / original_size_list must have one entry per frame")  # LISA.py:193
        offsets = [0]
        for e in seg_embeds:
            if e.dim() != 2 or e.shape[1] != OUT_CH:
                raise ValueError(f"seg_embeds entries must be (N,{OUT_CH})")
            offsets.append(offsets[-1] + e.shape[0])
        dev = frames.device
        self._ensure(dev, batch=B, prompts=offsets[-1])
        self._encode(frames, want_output=False)
        seg = torch.cat([e.to(device=dev, dtype=torch.float32) for e in seg_embeds], 0) if B else \
            torch.empty((0, OUT_CH), device=dev)
        res = self._decode_raw(None, B, seg, offsets, True, True, dev)
        out: Dict[str, object] = {"taxonomy": res["taxonomy"], "iou_left": res["iou_left"],
                                  "iou_right": res["iou_right"], "offsets": offsets,
                                  "lowres_left": res["lowres_left"], "lowres_right": res["lowres_right"]}
        for side in ("left", "right"):
            masks: List[torch.Tensor] = [None] * B
            logits: List[torch.Tensor] = [None] * B
            i = 0
            while i < B:   # one launch per run of frames that share (resize, original) sizes
                j = i + 1
                key = (tuple(int(v) for v in resize_list[i]), tuple(int(v) for v in original_size_list[i]))
                while j < B and (tuple(int(v) for v in resize_list[j]),
                                 tuple(int(v) for v in original_size_list[j])) == key:
                    j += 1
                low = res[f"lowres_{side}"][offsets[i]:offsets[j]]
                lg, mk = self._postprocess_raw(low, key[0], key[1], return_logits, True, threshold)
                if i == 0 and j == B:   # one size for the whole batch: also hand out the contiguous (P,H,W) buffer
                    out[f"masks_{side}_packed"] = mk
                for f in range(i, j):
                    a, b = offsets[f] - offsets[i], offsets[f + 1] - offsets[i]
                    masks[f] = mk[a:b]
                    if return_logits:
                        logits[f] = lg[a:b]
                i = j
            out[f"masks_{side}"] = masks
            if return_logits:
                out[f"logits_{side}"] = logits
        return out
