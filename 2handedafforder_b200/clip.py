"""Frame-sharded segmentation of a video clip over the GPUs of one box (BASELINE configs 3 and 5).

Frames are independent, so the clip is cut into contiguous shards - rank r owns frames
`[r*F/R, (r+1)*F/R)` - and every rank runs encode + both decoders + postprocess + threshold on its own shard in
micro-batches with replicated weights.  There is no collective on the data path; the ONE exchange step is the
gather of the thresholded masks (and the small taxonomy / IoU heads) to rank 0:

* wire format: 1 bit per pixel (`B200Sam.segment_bits`, numpy.packbits order) - 1.56 MB instead of 12.4 MB per 1080p
  frame with 3 prompts - in ONE flat message per micro-batch: `[bits | heads | per-frame prompt counts | checksums]`;
* zero staging: the postprocess kernel and the decoder heads write straight into the send slot, which is the buffer
  the gather reads (two slots, so the gather of micro-batch i overlaps the kernels of micro-batch i+1);
* rank 0 receives every round into one contiguous block, checks the per-mask 64-bit checksums the producers
  computed (`b200sam_checksum64`) against what arrived, moves the block to pinned host memory with ONE copy, and a
  consumer thread hands it to the caller - the launch thread never runs per-frame Python.

The reference has no counterpart (it evaluates one image at a time on one GPU, LISA.py:487-532); the per-frame
results are defined by `B200Sam.segment_bits`, and `tests/test_clip_gloo.py` checks that what rank 0 collects is
identical to a single-process run.

The module is transport-agnostic: `torch.distributed` with the NCCL backend on GPUs, gloo in the CPU tests (where
`segment_fn` is a stand-in; the product path always passes `B200Sam.segment_bits`, which has no CPU fallback).
"""
from __future__ import annotations

import queue
import threading
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import contextlib

import torch
import torch.distributed as dist

OUT_CH = 256


@contextlib.contextmanager
def _nvtx(name: str):
    """NVTX range per pipeline stage (nsys timelines, `ncu --nvtx`); a no-op without CUDA."""
    pushed = False
    try:
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_push(name)
            pushed = True
    except Exception:
        pushed = False
    try:
        yield
    finally:
        if pushed:
            torch.cuda.nvtx.range_pop()


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


def bits_row_bytes(width: int) -> int:
    return (int(width) + 7) // 8


def unpack_bits(bits: torch.Tensor, width: int) -> torch.Tensor:
    """(..., H, ceil(W/8)) packed u8 -> (..., H, W) u8 {0,1} (numpy.packbits(axis=-1) bit order)."""
    shifts = torch.arange(7, -1, -1, device=bits.device, dtype=torch.uint8)
    full = (bits.unsqueeze(-1) >> shifts) & 1
    return full.reshape(*bits.shape[:-1], bits.shape[-1] * 8)[..., :width].contiguous()


def _up(n: int, a: int = 16) -> int:
    return (n + a - 1) // a * a


class SlotLayout:
    """Byte layout of one micro-batch message:
    `[bits (2,cap,H,Wb) u8 | heads (3,cap,4) f32 | counts (mb) i64 | sums (2*cap) i64]`, every part 16-byte aligned.
    heads = [taxonomy, iou_left, iou_right]; counts = prompts per frame; sums = checksum of every (side, prompt) mask."""

    def __init__(self, micro_batch: int, max_prompts_per_frame: int, hw: Tuple[int, int]):
        self.mb, self.cap = int(micro_batch), int(micro_batch) * int(max_prompts_per_frame)
        self.H, self.W = int(hw[0]), int(hw[1])
        self.Wb = bits_row_bytes(self.W)
        self.mask_bytes = self.H * self.Wb
        self.o_heads = _up(2 * self.cap * self.mask_bytes)
        self.o_counts = self.o_heads + _up(3 * self.cap * 4 * 4)
        self.o_sums = self.o_counts + _up(8 * self.mb)
        self.nbytes = self.o_sums + _up(8 * 2 * self.cap)

    def views(self, buf: torch.Tensor):
        """Typed views into a flat (nbytes,) u8 buffer: bits (2,cap,H,Wb), heads (3,cap,4), counts (mb,), sums (2*cap,)."""
        cap = self.cap
        bits = buf[: 2 * cap * self.mask_bytes].view(2, cap, self.H, self.Wb)
        heads = buf[self.o_heads: self.o_heads + 3 * cap * 16].view(torch.float32).view(3, cap, 4)
        counts = buf[self.o_counts: self.o_counts + 8 * self.mb].view(torch.int64)
        sums = buf[self.o_sums: self.o_sums + 16 * cap].view(torch.int64)
        return bits, heads, counts, sums


@dataclass
class ClipResult:
    """What rank `dst` holds after `segment_clip` (other ranks get an empty one).  Index = global frame number; all
    tensors live in host memory, masks stay in the packed wire format (`masks(f, side)` unpacks one frame)."""
    bits_left: List[Optional[torch.Tensor]] = field(default_factory=list)     # (N_i,H,ceil(W/8)) u8 per frame
    bits_right: List[Optional[torch.Tensor]] = field(default_factory=list)
    taxonomy: List[Optional[torch.Tensor]] = field(default_factory=list)      # (N_i,4) f32
    iou_left: List[Optional[torch.Tensor]] = field(default_factory=list)      # (N_i,4) f32
    iou_right: List[Optional[torch.Tensor]] = field(default_factory=list)
    frames_done: int = 0
    width: int = 0
    masks_checked: int = 0           # (side, prompt) masks whose checksum was compared after the gather
    checksum_mismatches: int = 0     # ... and how many differed (0 = what arrived is what the producers wrote)

    def masks(self, frame: int, side: str = "left") -> torch.Tensor:
        """(N_i,H,W) u8 {0,1} thresholded masks of one frame."""
        b = (self.bits_left if side == "left" else self.bits_right)[frame]
        return unpack_bits(b, self.width)


class _Consumer(threading.Thread):
    """Rank dst: drains rounds from pinned host blocks in arrival order - parses the per-frame prompt counts, calls the
    caller's sinks, fills the result lists - so that the thread that launches kernels never waits for a copy and never
    runs per-frame Python.  `free` hands the host blocks back to the producer side."""

    def __init__(self, lay: SlotLayout, n_frames: int, world: int, micro_batch: int, res: ClipResult, keep: bool,
                 sink, batch_sink, host: List[torch.Tensor]):
        super().__init__(daemon=True, name="b200sam-clip-consumer")
        self.lay, self.n_frames, self.world, self.mb = lay, n_frames, world, micro_batch
        self.res, self.keep, self.sink, self.batch_sink, self.host = res, keep, sink, batch_sink, host
        self.todo: "queue.Queue" = queue.Queue()
        self.free: "queue.Queue" = queue.Queue()
        for i in range(len(host)):
            self.free.put(i)
        self.error: Optional[BaseException] = None

    def run(self) -> None:
        while True:
            item = self.todo.get()
            if item is None:
                return
            rnd, hidx, ev = item
            try:
                if self.error is None:
                    if ev is not None:
                        ev.synchronize()          # the D2H of this round (releases the GIL while it waits)
                    self._drain(rnd, self.host[hidx])
            except BaseException as e:           # surfaced by segment_clip after the join
                self.error = e
            finally:
                self.free.put(hidx)

    def _drain(self, rnd: int, block: torch.Tensor) -> None:
        lay, res = self.lay, self.res
        for src in range(self.world):
            a, b = round_frames(self.n_frames, src, self.world, self.mb, rnd)
            if b <= a:
                continue
            bits, heads, counts, _ = lay.views(block[src])
            cnt = counts[: b - a].tolist()
            P = sum(cnt)
            if self.batch_sink is not None:
                self.batch_sink(src, a, cnt, bits[:, :P], heads[:, :P])
            off = 0
            for f, n in zip(range(a, b), cnt):
                if self.sink is not None:
                    self.sink(f, bits[0, off:off + n], bits[1, off:off + n], heads[0, off:off + n],
                              heads[1, off:off + n], heads[2, off:off + n])
                if self.keep:
                    res.bits_left[f], res.bits_right[f] = bits[0, off:off + n].clone(), bits[1, off:off + n].clone()
                    res.taxonomy[f], res.iou_left[f] = heads[0, off:off + n].clone(), heads[1, off:off + n].clone()
                    res.iou_right[f] = heads[2, off:off + n].clone()
                off += n
            res.frames_done += b - a


def segment_clip(segment_fn: Callable[..., Dict[str, object]],
                 load_frames: Callable[[int, int], Tuple[torch.Tensor, Sequence[torch.Tensor]]],
                 n_frames: int, resize: Tuple[int, int], original: Tuple[int, int], *,
                 micro_batch: int = 8, max_prompts_per_frame: int = 3, threshold: float = 0.0,
                 device: Optional[torch.device] = None, group=None, dst: int = 0,
                 sink: Optional[Callable[..., None]] = None, batch_sink: Optional[Callable[..., None]] = None,
                 keep: bool = True, prefetch: bool = False,
                 checksum_fn: Optional[Callable[..., torch.Tensor]] = None, host_blocks: int = 3) -> ClipResult:
    """Segment frames [0, n_frames) of a clip on all ranks of `group`; rank `dst` collects the results on the host.

    segment_fn: `B200Sam.segment_bits` (frames, seg_embeds, resize, original, threshold=, bits_out=, heads_out=):
        writes the packed masks / heads of the micro-batch into the given views of the send slot.
    load_frames(lo, hi): this rank's input for global frames [lo, hi): `(frames (hi-lo,3,1024,1024), [seg (N_i,256)])`
        already on `device` (or pinned host memory the caller copies itself).
    checksum_fn: `B200Sam.checksum64` (u8 (n, ...) -> int64 (n,), optional out=).  When given, every producer ships
        the checksum of each (side, prompt) mask and rank dst recomputes it on what the gather delivered;
        `ClipResult.masks_checked / checksum_mismatches` report the outcome.
    sink(frame, bits_left, bits_right, taxonomy, iou_left, iou_right): called on rank dst once per frame, in arrival
        order, from the consumer thread, with HOST views (valid only during the call).
    batch_sink(src_rank, first_frame, counts, bits (2,P,H,Wb), heads (3,P,4)): same, once per (round, source rank) -
        the streaming consumer of config 5 (keep=False so nothing accumulates).
    prefetch: on GPUs, call load_frames for round r+1 on an upload stream while round r computes (for loaders that
        copy from pinned host memory with non_blocking=True: the H2D copy leaves the critical path).
    Every frame of the clip shares (resize, original), as a video does.
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if micro_batch <= 0 or max_prompts_per_frame <= 0:
        raise ValueError("micro_batch and max_prompts_per_frame must be positive")
    lay = SlotLayout(micro_batch, max_prompts_per_frame, original)
    rounds = n_rounds(n_frames, world, micro_batch)
    res = ClipResult(width=lay.W)
    is_dst = rank == dst
    if is_dst and keep:
        for lst in (res.bits_left, res.bits_right, res.taxonomy, res.iou_left, res.iou_right):
            lst.extend([None] * n_frames)
    if rounds == 0:
        return res
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    on_gpu = device.type == "cuda"
    u8 = dict(dtype=torch.uint8, device=device)
    send = [torch.zeros((lay.nbytes,), **u8) for _ in range(2)]
    send_work: List[Optional[object]] = [None, None]
    recv = [torch.zeros((world, lay.nbytes), **u8) for _ in range(2)] if (is_dst and world > 1) else None
    main = torch.cuda.current_stream(device) if on_gpu else None
    side = torch.cuda.Stream(device=device) if on_gpu else None          # gather completion, verification, D2H
    d2h_done: List[Optional[object]] = [None, None]                       # recv[k] may be overwritten after this
    consumer = None
    mism = chk_all = None
    if is_dst:
        host = [torch.empty((world, lay.nbytes), dtype=torch.uint8) for _ in range(max(2, host_blocks))]
        if on_gpu:
            host = [h.pin_memory() for h in host]
        consumer = _Consumer(lay, n_frames, world, micro_batch, res, keep, sink, batch_sink, host)
        consumer.start()
        if checksum_fn is not None:
            mism = torch.zeros((), dtype=torch.int64, device=device)
            chk_all = torch.zeros((world, 2 * lay.cap), dtype=torch.int64, device=device)

    up = torch.cuda.Stream(device=device) if (prefetch and on_gpu) else None

    def fetch(rnd: int):
        """This rank's input of round `rnd` (None when its shard is exhausted), loaded on the upload stream."""
        a, b = round_frames(n_frames, rank, world, micro_batch, rnd)
        if rnd >= rounds or b <= a:
            return None
        if up is None:
            return load_frames(a, b) + (None,)
        with torch.cuda.stream(up), _nvtx("clip.load_frames"):
            frames, segs = load_frames(a, b)
            ev = torch.cuda.Event()
            ev.record(up)
        return frames, segs, ev

    def deliver(rnd: int, block_dev: torch.Tensor) -> None:
        """Rank dst: verify what arrived, move the round's block to a pinned host block with ONE copy (side stream)
        and queue it for the consumer thread."""
        hidx = consumer.free.get()                   # back-pressure: at most `host_blocks` rounds in flight
        hb = consumer.host[hidx]
        ctx = torch.cuda.stream(side) if on_gpu else _NullCtx()
        with ctx:
            if world > 1:
                send_work[rnd % 2].wait()            # NCCL: orders the side stream behind the gather, host does not block
            elif on_gpu:
                side.wait_stream(main)
            if checksum_fn is not None:
                for src in range(world):
                    bits, _, _, _ = lay.views(block_dev[src])
                    checksum_fn(bits.view(2 * lay.cap, lay.H, lay.Wb), out=chk_all[src])
                sums = block_dev[:, lay.o_sums: lay.o_sums + 16 * lay.cap].view(torch.int64)
                mism.add_((chk_all != sums).sum())
                res.masks_checked += world * 2 * lay.cap
            hb.copy_(block_dev, non_blocking=True)
            ev = None
            if on_gpu:
                ev = torch.cuda.Event()
                ev.record(side)
                d2h_done[rnd % 2] = ev
        consumer.todo.put((rnd, hidx, ev))

    try:
        nxt = fetch(0)
        for rnd in range(rounds):
            k = rnd % 2
            slot = send[k]
            bits, heads, counts_t, sums = lay.views(slot)
            if send_work[k] is not None:
                send_work[k].wait()                  # the gather that last read this slot (round rnd-2) is done
                send_work[k] = None
            if on_gpu and d2h_done[k] is not None and world == 1:
                main.wait_event(d2h_done[k])         # world 1: the D2H reads the send slot itself
            a, b = round_frames(n_frames, rank, world, micro_batch, rnd)
            cur, nxt = nxt, (fetch(rnd + 1) if up is not None else None)
            counts = [0] * micro_batch
            if b > a:
                if cur is None:
                    cur = fetch(rnd)
                frames, segs, ev = cur
                if ev is not None:                   # loaded on the upload stream: order it before the kernels
                    main.wait_event(ev)
                    for t in [frames, *segs]:
                        t.record_stream(main)
                if len(segs) != b - a or frames.shape[0] != b - a:
                    raise ValueError("load_frames returned the wrong number of frames")
                counts[: b - a] = [int(s.shape[0]) for s in segs]
                if max(counts) > max_prompts_per_frame:
                    raise ValueError(f"a frame has more than max_prompts_per_frame={max_prompts_per_frame} prompts")
                with _nvtx("clip.segment"):
                    segment_fn(frames, segs, resize, original, threshold=threshold, bits_out=bits, heads_out=heads)
            counts_t.copy_(torch.tensor(counts, dtype=torch.int64), non_blocking=True)
            if checksum_fn is not None:
                checksum_fn(bits.view(2 * lay.cap, lay.H, lay.Wb), out=sums)
            if world > 1:
                if is_dst and on_gpu and d2h_done[k] is not None:
                    main.wait_event(d2h_done[k])     # recv[k] was drained to the host (round rnd-2)
                # the one exchange step: ONE gather per round, asynchronous so the next micro-batch's kernels
                # overlap the NVLink transfer
                dest = [recv[k][s] for s in range(world)] if is_dst else None
                with _nvtx("clip.gather"):
                    send_work[k] = dist.gather(slot, dest, dst=dst, group=group, async_op=True)
                if is_dst:
                    deliver(rnd, recv[k])
            elif is_dst:
                deliver(rnd, slot.view(1, -1))
    finally:
        if consumer is not None:
            consumer.todo.put(None)
            consumer.join()
    for w in send_work:
        if w is not None:
            w.wait()
    if on_gpu:
        main.wait_stream(side)
    if consumer is not None and consumer.error is not None:
        raise consumer.error
    if mism is not None:
        res.checksum_mismatches = int(mism.item())
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
    memory and the result of micro-batch k-1 is downloaded on two copy streams while the kernels of micro-batch k run;
    two input and two output slots, event-ordered, no host synchronisation inside the loop.  The result travels in the
    1-bit wire format as ONE flat message per micro-batch (`SlotLayout`: bits | heads | counts | checksums) that the
    kernels write in place - one D2H copy per step.

        pipe = HostPipeline(sam.segment_bits, batch, prompts_per_frame, resize, original, device)
        for k in range(n): pipe.submit(host_frames_k, host_seg_k)   # pinned (B,3,1024,1024) / (B*NP,256)
        pipe.flush()                                                 # result of submit k: pipe.result(k % 2)
    """

    def __init__(self, segment_fn, batch: int, prompts_per_frame: int, resize, original, device: torch.device,
                 frame_dtype: torch.dtype = torch.bfloat16, threshold: float = 0.0):
        self.segment_fn, self.B, self.NP = segment_fn, int(batch), int(prompts_per_frame)
        self.resize, self.original, self.device, self.threshold = resize, original, device, threshold
        self.lay = SlotLayout(self.B, self.NP, original)
        P = self.B * self.NP
        self.up = torch.cuda.Stream(device=device)       # H2D and D2H run on their own streams (two copy engines)
        self.down = torch.cuda.Stream(device=device)
        self.d_frames = [torch.empty((self.B, 3, 1024, 1024), dtype=frame_dtype, device=device) for _ in range(2)]
        self.d_seg = [torch.empty((P, OUT_CH), dtype=torch.float32, device=device) for _ in range(2)]
        self.d_out = [torch.zeros((self.lay.nbytes,), dtype=torch.uint8, device=device) for _ in range(2)]
        self.host_out = [torch.zeros((self.lay.nbytes,), dtype=torch.uint8).pin_memory() for _ in range(2)]
        self.up_done = [torch.cuda.Event() for _ in range(2)]        # upload of slot finished
        self.in_free = [torch.cuda.Event() for _ in range(2)]        # kernels that read the input slot finished
        self.down_done = [torch.cuda.Event() for _ in range(2)]      # download into host_out[slot] finished
        self.k = 0
        self.h2d_bytes = self.d_frames[0].numel() * self.d_frames[0].element_size() + self.d_seg[0].numel() * 4
        self.d2h_bytes = self.lay.nbytes

    def submit(self, host_frames: torch.Tensor, host_seg: torch.Tensor) -> int:
        s = self.k % 2
        main = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self.up), _nvtx("pipeline.h2d"):
            if self.k >= 2:
                self.up.wait_event(self.in_free[s])                  # the step that last read this slot is done
            self.d_frames[s].copy_(host_frames, non_blocking=True)
            self.d_seg[s].copy_(host_seg, non_blocking=True)
            self.up_done[s].record(self.up)
        main.wait_event(self.up_done[s])
        if self.k >= 2:
            main.wait_event(self.down_done[s])                       # d_out[s] was downloaded (step k-2)
        segs = [self.d_seg[s][i * self.NP:(i + 1) * self.NP] for i in range(self.B)]
        bits, heads, counts, _ = self.lay.views(self.d_out[s])
        with _nvtx("pipeline.segment"):
            self.segment_fn(self.d_frames[s], segs, self.resize, self.original, threshold=self.threshold, bits_out=bits,
                            heads_out=heads)
        if self.k < 2:
            counts.fill_(self.NP)
        self.in_free[s].record(main)
        with torch.cuda.stream(self.down), _nvtx("pipeline.d2h"):
            self.down.wait_event(self.in_free[s])
            self.host_out[s].copy_(self.d_out[s], non_blocking=True)
            self.down_done[s].record(self.down)
        self.k += 1
        return s

    def result(self, slot: int) -> Dict[str, torch.Tensor]:
        """Host views of the message in `host_out[slot]` (valid once `down_done[slot]` has completed)."""
        bits, heads, counts, _ = self.lay.views(self.host_out[slot])
        return {"bits": bits, "heads": heads, "counts": counts}

    def flush(self) -> None:
        self.up.synchronize()
        self.down.synchronize()
        torch.cuda.current_stream(self.device).synchronize()
