"""Frame-sharded segmentation of a video clip over the GPUs of one box (BASELINE configs 3 and 5).

Frames are independent, so the clip is cut into contiguous shards - rank r owns frames
`[r*F/R, (r+1)*F/R)` - and every rank runs encode + both decoders + postprocess + threshold on its own shard in
micro-batches with replicated weights.  There is no collective on the data path; the ONE exchange step is the
gather of the thresholded u8 masks (and the small taxonomy / IoU heads) to rank 0.  The gather of micro-batch i is
issued asynchronously from a double-buffered send slot, so it overlaps the kernels of micro-batch i+1.

The reference has no counterpart (it evaluates one image at a time on one GPU, LISA.py:487-532); the per-frame
results are defined by `B200Sam.segment`, and `tests/test_clip_gloo.py` checks that what rank 0 collects is
identical to a single-process run.

The module is transport-agnostic: `torch.distributed` with the NCCL backend on GPUs, gloo in the CPU tests (where
`segment_fn` is a stand-in; the product path always passes `B200Sam.segment`, which has no CPU fallback).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

OUT_CH = 256


def shard_range(n_frames: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard of rank `rank`: [rank*F/R, (rank+1)*F/R) with integer floor - sizes differ by at most 1."""
    if n_frames < 0 or world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad shard request: n_frames={n_frames} rank={rank} world={world}")
    return (rank * n_frames) // world, ((rank + 1) * n_frames) // world


def n_rounds(n_frames: int, world: int, micro_batch: int) -> int:
    """Micro-batch rounds every rank takes part in (the largest shard decides; short shards send padding)."""
    longest = max(b - a for a, b in (shard_range(n_frames, r, world) for r in range(world))) if world else 0
    return -(-longest // micro_batch)


def round_frames(n_frames: int, rank: int, world: int, micro_batch: int, rnd: int) -> Tuple[int, int]:
    """Global frame interval rank `rank` processes in round `rnd` (empty once its shard is exhausted)."""
    lo, hi = shard_range(n_frames, rank, world)
    a = min(lo + rnd * micro_batch, hi)
    return a, min(a + micro_batch, hi)


@dataclass
class ClipResult:
    """What rank 0 holds after `segment_clip` (other ranks get an empty one).  Index = global frame number."""
    masks_left: List[Optional[torch.Tensor]] = field(default_factory=list)    # (N_i,H,W) u8 per frame
    masks_right: List[Optional[torch.Tensor]] = field(default_factory=list)
    taxonomy: List[Optional[torch.Tensor]] = field(default_factory=list)      # (N_i,4) f32
    iou_left: List[Optional[torch.Tensor]] = field(default_factory=list)      # (N_i,4) f32
    iou_right: List[Optional[torch.Tensor]] = field(default_factory=list)
    frames_done: int = 0


class _Slot:
    """One send slot: packed masks (2, mb*max_prompts, H, W) u8, heads (3, mb*max_prompts, 4) f32, counts (mb,) i32."""

    def __init__(self, mb: int, max_prompts: int, hw: Tuple[int, int], device: torch.device):
        cap = mb * max_prompts
        self.masks = torch.empty((2, cap, hw[0], hw[1]), dtype=torch.uint8, device=device)
        self.heads = torch.empty((3, cap, 4), dtype=torch.float32, device=device)
        self.counts = torch.zeros((mb,), dtype=torch.int32, device=device)
        self.works: list = []

    def wait(self) -> None:
        for w in self.works:
            w.wait()
        self.works = []


def segment_clip(segment_fn: Callable[..., Dict[str, object]],
                 load_frames: Callable[[int, int], Tuple[torch.Tensor, Sequence[torch.Tensor]]],
                 n_frames: int, resize: Tuple[int, int], original: Tuple[int, int], *,
                 micro_batch: int = 8, max_prompts_per_frame: int = 3, threshold: float = 0.0,
                 device: Optional[torch.device] = None, group=None, dst: int = 0,
                 sink: Optional[Callable[[int, torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor,
                                          torch.Tensor], None]] = None,
                 keep: bool = True, prefetch: bool = False) -> ClipResult:
    """Segment frames [0, n_frames) of a clip on all ranks of `group`; rank `dst` collects the results.

    segment_fn: `B200Sam.segment` (frames, seg_embeds, resize_list, original_size_list, threshold=...) -> dict.
    load_frames(lo, hi): this rank's input for global frames [lo, hi): `(frames (hi-lo,3,1024,1024), [seg (N_i,256)])`
        already on `device` (or pinned host memory the caller copies itself).
    sink(frame, masks_left, masks_right, taxonomy, iou_left, iou_right): called on rank `dst` once per frame, in
        arrival order, with views into the receive buffers (valid only during the call) - streaming consumers
        (config 5) pass keep=False so nothing accumulates.
    prefetch: on GPUs, call load_frames for round r+1 on an upload stream while round r computes (for loaders that
        copy from pinned host memory with non_blocking=True: the H2D copy leaves the critical path).
    Every frame of the clip shares (resize, original), as a video does.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if micro_batch <= 0 or max_prompts_per_frame <= 0:
        raise ValueError("micro_batch and max_prompts_per_frame must be positive")
    H, W = int(original[0]), int(original[1])
    rounds = n_rounds(n_frames, world, micro_batch)
    res = ClipResult()
    is_dst = rank == dst
    if is_dst and keep:
        for lst in (res.masks_left, res.masks_right, res.taxonomy, res.iou_left, res.iou_right):
            lst.extend([None] * n_frames)
    if rounds == 0:
        return res
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    slots = [_Slot(micro_batch, max_prompts_per_frame, (H, W), device) for _ in range(2)]
    recv = None
    if is_dst and world > 1:   # double-buffered like the send side: round i is consumed while round i+1 arrives
        recv = [[_Slot(micro_batch, max_prompts_per_frame, (H, W), device) for _ in range(world)] for _ in range(2)]

    on_gpu = device.type == "cuda"
    side = torch.cuda.Stream(device=device) if (on_gpu and world > 1) else None

    def consume(rnd: int) -> None:
        """Rank dst: hand the frames of round `rnd` (from every source rank) to the sink / result lists.  On GPUs
        this runs on a side stream, so the host only waits for the gather of round `rnd` - never for the kernels of
        the micro-batch already queued behind it on the main stream."""
        ctx = torch.cuda.stream(side) if side is not None else _NullCtx()
        with ctx:
            if world > 1:
                slots[rnd % 2].wait()                 # gather of round rnd complete (stream-ordered under NCCL)
            for src in range(world):
                slot = slots[rnd % 2] if world == 1 else recv[rnd % 2][src]
                a, b = round_frames(n_frames, src, world, micro_batch, rnd)
                counts = slot.counts[:b - a].tolist()
                off = 0
                for f, n in zip(range(a, b), counts):
                    ml, mr = slot.masks[0, off:off + n], slot.masks[1, off:off + n]
                    tx, il, ir = slot.heads[0, off:off + n], slot.heads[1, off:off + n], slot.heads[2, off:off + n]
                    if sink is not None:
                        sink(f, ml, mr, tx, il, ir)
                    if keep:
                        res.masks_left[f], res.masks_right[f] = ml.clone(), mr.clone()
                        res.taxonomy[f], res.iou_left[f], res.iou_right[f] = tx.clone(), il.clone(), ir.clone()
                    res.frames_done += 1
                    off += n

    up = torch.cuda.Stream(device=device) if (prefetch and on_gpu) else None

    def fetch(rnd: int):
        """This rank's input of round `rnd` (None when its shard is exhausted), loaded on the upload stream."""
        a, b = round_frames(n_frames, rank, world, micro_batch, rnd)
        if rnd >= rounds or b <= a:
            return None
        if up is None:
            return load_frames(a, b) + (None,)
        with torch.cuda.stream(up):
            frames, segs = load_frames(a, b)
            ev = torch.cuda.Event()
            ev.record(up)
        return frames, segs, ev

    nxt = fetch(0)
    for rnd in range(rounds):
        slot = slots[rnd % 2]
        if not is_dst:
            slot.wait()                               # the gather that last read this slot (round rnd-2) is done
        a, b = round_frames(n_frames, rank, world, micro_batch, rnd)
        slot.counts.zero_()
        cur, nxt = nxt, (fetch(rnd + 1) if up is not None else None)
        if b > a:
            if cur is None:
                cur = fetch(rnd)
            frames, segs, ev = cur
            if ev is not None:                        # loaded on the upload stream: order it before the kernels
                main = torch.cuda.current_stream(device)
                main.wait_event(ev)
                for t in [frames, *segs]:
                    t.record_stream(main)
            if len(segs) != b - a or frames.shape[0] != b - a:
                raise ValueError("load_frames returned the wrong number of frames")
            counts = [int(s.shape[0]) for s in segs]
            if max(counts, default=0) > max_prompts_per_frame:
                raise ValueError(f"a frame has more than max_prompts_per_frame={max_prompts_per_frame} prompts")
            out = segment_fn(frames, segs, [resize] * (b - a), [original] * (b - a), threshold=threshold)
            P = sum(counts)
            if P:
                slot.masks[0, :P].copy_(out["masks_left_packed"])
                slot.masks[1, :P].copy_(out["masks_right_packed"])
                slot.heads[0, :P].copy_(out["taxonomy"])
                slot.heads[1, :P].copy_(out["iou_left"])
                slot.heads[2, :P].copy_(out["iou_right"])
            slot.counts[:b - a].copy_(torch.tensor(counts, dtype=torch.int32))
        if world > 1:
            if side is not None:                      # recv[rnd % 2] must have been drained (round rnd-2's consume)
                torch.cuda.current_stream(device).wait_stream(side)
            # the one exchange step: three gathers per round (u8 masks, f32 heads, i32 counts), asynchronous so the
            # next micro-batch's kernels overlap the NVLink transfer
            for name in ("masks", "heads", "counts"):
                t = getattr(slot, name)
                dest = [getattr(recv[rnd % 2][s], name) for s in range(world)] if is_dst else None
                slot.works.append(dist.gather(t, dest, dst=dst, group=group, async_op=True))
            if is_dst and rnd >= 1:
                consume(rnd - 1)                      # arrived while round rnd was being computed
        elif is_dst:
            consume(rnd)
    if world > 1:
        if is_dst:
            consume(rounds - 1)
        for s in slots:
            s.wait()
        if side is not None:
            torch.cuda.current_stream(device).wait_stream(side)
    return res


class _NullCtx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class HostPipeline:
    """Host-to-host segmentation of a stream of micro-batches with the PCIe copies off the critical path.

    The reference moves one image at a time (`.cuda()` before `evaluate`, `.cpu().numpy()` after,
    inference.py:247-297) and so serialises copy - compute - copy.  Here micro-batch k+1 is uploaded from pinned host
    memory and the masks of micro-batch k-1 are downloaded on two copy streams while the kernels of micro-batch k run;
    two input and two output slots, event-ordered, no host synchronisation inside the loop.

        pipe = HostPipeline(sam.segment, batch, prompts_per_frame, resize, original, device)
        for k in range(n): pipe.submit(host_frames_k, host_seg_k)   # pinned (B,3,1024,1024) / (B*NP,256)
        pipe.flush()                                                 # results of submit k are in pipe.host_out[k % 2]
    """

    def __init__(self, segment_fn, batch: int, prompts_per_frame: int, resize, original, device: torch.device,
                 frame_dtype: torch.dtype = torch.bfloat16, threshold: float = 0.0):
        self.segment_fn, self.B, self.NP = segment_fn, int(batch), int(prompts_per_frame)
        self.resize, self.original, self.device, self.threshold = resize, original, device, threshold
        P, (H, W) = self.B * self.NP, (int(original[0]), int(original[1]))
        self.up = torch.cuda.Stream(device=device)       # H2D and D2H run on their own streams (two copy engines)
        self.down = torch.cuda.Stream(device=device)
        self.d_frames = [torch.empty((self.B, 3, 1024, 1024), dtype=frame_dtype, device=device) for _ in range(2)]
        self.d_seg = [torch.empty((P, OUT_CH), dtype=torch.float32, device=device) for _ in range(2)]
        self.host_out = [{"masks": torch.empty((2, P, H, W), dtype=torch.uint8).pin_memory(),
                          "heads": torch.empty((3, P, 4), dtype=torch.float32).pin_memory()} for _ in range(2)]
        self.up_done = [torch.cuda.Event() for _ in range(2)]        # upload of slot finished
        self.in_free = [torch.cuda.Event() for _ in range(2)]        # kernels that read the input slot finished
        self.down_done = [torch.cuda.Event() for _ in range(2)]      # download into host_out[slot] finished
        self.k = 0
        self.h2d_bytes = self.d_frames[0].numel() * self.d_frames[0].element_size() + self.d_seg[0].numel() * 4
        self.d2h_bytes = self.host_out[0]["masks"].numel() + self.host_out[0]["heads"].numel() * 4

    def submit(self, host_frames: torch.Tensor, host_seg: torch.Tensor) -> int:
        s = self.k % 2
        main = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self.up):
            if self.k >= 2:
                self.up.wait_event(self.in_free[s])                  # the step that last read this slot is done
            self.d_frames[s].copy_(host_frames, non_blocking=True)
            self.d_seg[s].copy_(host_seg, non_blocking=True)
            self.up_done[s].record(self.up)
        main.wait_event(self.up_done[s])
        segs = [self.d_seg[s][i * self.NP:(i + 1) * self.NP] for i in range(self.B)]
        out = self.segment_fn(self.d_frames[s], segs, [self.resize] * self.B, [self.original] * self.B,
                              threshold=self.threshold)
        self.in_free[s].record(main)
        with torch.cuda.stream(self.down):
            self.down.wait_event(self.in_free[s])
            ho = self.host_out[s]
            ho["masks"][0].copy_(out["masks_left_packed"], non_blocking=True)
            ho["masks"][1].copy_(out["masks_right_packed"], non_blocking=True)
            ho["heads"][0].copy_(out["taxonomy"], non_blocking=True)
            ho["heads"][1].copy_(out["iou_left"], non_blocking=True)
            ho["heads"][2].copy_(out["iou_right"], non_blocking=True)
            self.down_done[s].record(self.down)
            for t in (out["masks_left_packed"], out["masks_right_packed"], out["taxonomy"], out["iou_left"],
                      out["iou_right"]):
                t.record_stream(self.down)                           # allocator: do not recycle before the download
        self.k += 1
        return s

    def flush(self) -> None:
        self.up.synchronize()
        self.down.synchronize()
        torch.cuda.current_stream(self.device).synchronize()
