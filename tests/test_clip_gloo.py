"""world_size-2 (and 3) gloo tests of the frame-sharded clip path (2handedafforder_b200/clip.py).

The exchange logic (shard ranges, the flat 1-bit message layout, double-buffered asynchronous gather, checksum
verification, ragged shards, variable prompt counts, the consumer thread with its per-frame and per-round sinks) is
host code and is exercised here on CPU with a deterministic stand-in for `B200Sam.segment_bits`; the
real kernels are covered by the `-m gpu` tests and by `bench.py --gpus N`.  What rank 0 collects must be identical
to a single-process run over the same clip.
"""
import importlib
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import aux_oracle as A
from tests import gpu_util as G

clip = importlib.import_module(G.PKG + ".clip")

HW = (6, 10)


def _frame(i):
    g = torch.Generator().manual_seed(1000 + i)
    return torch.randn((3, 4, 4), generator=g)


def _n_prompts(i):
    return (i * 7 + 1) % 4          # 0..3 prompts: includes frames with no [SEG] at all (chat.py:227)


def _seg(i):
    g = torch.Generator().manual_seed(5000 + i)
    return torch.randn((_n_prompts(i), 256), generator=g)


def _load(lo, hi):
    return torch.stack([_frame(i) for i in range(lo, hi)]), [_seg(i) for i in range(lo, hi)]


def _fake_masks(frames, segs, H, W, threshold):
    """Deterministic stand-in for the arithmetic of B200Sam.segment_bits: u8 masks + heads per prompt."""
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    ml, mr, tx, il, ir = [], [], [], [], []
    for f, s in zip(frames, segs):
        for p in range(s.shape[0]):
            v = f.sum() + s[p, :8].sum()
            ml.append((torch.sin(yy * v + xx) > threshold).to(torch.uint8))
            mr.append((torch.cos(xx * v - yy) > threshold).to(torch.uint8))
            tx.append(torch.softmax(s[p, :4], 0))
            il.append(s[p, 4:8] * v)
            ir.append(s[p, 8:12] - v)
    st = lambda l, shape: torch.stack(l) if l else torch.zeros((0, *shape))
    return (st(ml, (H, W)).to(torch.uint8), st(mr, (H, W)).to(torch.uint8), st(tx, (4,)), st(il, (4,)), st(ir, (4,)))


def _fake_segment_bits(frames, segs, resize, original, threshold=0.0, bits_out=None, heads_out=None):
    """Stand-in with the output contract of B200Sam.segment_bits: packs into the caller's send-slot views."""
    H, W = original
    ml, mr, tx, il, ir = _fake_masks(frames, segs, H, W, threshold)
    P = ml.shape[0]
    if P:
        bits_out[0, :P] = torch.from_numpy(A.pack_bits(ml.numpy()))
        bits_out[1, :P] = torch.from_numpy(A.pack_bits(mr.numpy()))
        heads_out[0, :P], heads_out[1, :P], heads_out[2, :P] = tx, il, ir
    return {}


def _checksum(data, out=None):
    """CPU stand-in for B200Sam.checksum64: the oracle's restatement of b200sam_checksum64."""
    v = torch.from_numpy(A.checksum64(data.numpy()))
    if out is None:
        return v
    out[: v.numel()] = v
    return out


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_frames, mb, path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        seen, batches = [], []
        res = clip.segment_clip(_fake_segment_bits, _load, n_frames, (4, 4), HW, micro_batch=mb,
                                max_prompts_per_frame=3, device=torch.device("cpu"), checksum_fn=_checksum,
                                sink=lambda f, *a: seen.append((f, int(clip.unpack_bits(a[0], HW[1]).sum()),
                                                                a[0].shape[0])),
                                batch_sink=lambda src, lo, cnt, bits, heads: batches.append(
                                    (src, lo, list(cnt), bits.shape[1], heads.shape[1])))
        if rank == 0:
            torch.save({"res": res, "seen": seen, "batches": batches}, path)
        else:
            assert res.frames_done == 0 and not res.bits_left
    finally:
        dist.destroy_process_group()


def _single(n_frames, mb):
    return clip.segment_clip(_fake_segment_bits, _load, n_frames, (4, 4), HW, micro_batch=mb, max_prompts_per_frame=3,
                             device=torch.device("cpu"), checksum_fn=_checksum)


def test_shard_ranges_cover_the_clip_exactly():
    for n in (0, 1, 7, 256, 2048, 2051):
        for world in (1, 2, 3, 8):
            spans = [clip.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
            for mb in (1, 8):
                rounds = clip.n_rounds(n, world, mb)
                got = []
                for r in range(world):
                    for k in range(rounds):
                        a, b = clip.round_frames(n, r, world, mb, k)
                        got.extend(range(a, b))
                    assert clip.round_frames(n, r, world, mb, rounds) [0] == spans[r][1]
                assert sorted(got) == list(range(n))
    with pytest.raises(ValueError):
        clip.shard_range(4, 2, 2)


def test_single_process_clip_matches_direct_calls():
    res = _single(5, 2)
    assert res.frames_done == 5
    assert res.checksum_mismatches == 0 and res.masks_checked > 0
    for i in range(5):
        ml, mr, tx, il, ir = _fake_masks(_frame(i)[None], [_seg(i)], *HW, 0.0)
        assert torch.equal(res.masks(i, "left"), ml) and torch.equal(res.masks(i, "right"), mr)
        assert torch.equal(res.iou_right[i], ir) and torch.equal(res.taxonomy[i], tx)
        assert res.bits_left[i].shape == (_n_prompts(i), HW[0], (HW[1] + 7) // 8)


@pytest.mark.parametrize("world,n_frames,mb", [(2, 13, 4), (2, 8, 8), (3, 7, 2), (2, 1, 4)])
def test_gathered_clip_is_identical_to_one_rank(world, n_frames, mb, tmp_path):
    path = str(tmp_path / "rank0.pt")
    mp.spawn(_worker, args=(world, _free_port(), n_frames, mb, path), nprocs=world, join=True)
    blob = torch.load(path, weights_only=False)
    got, ref = blob["res"], _single(n_frames, mb)
    assert got.frames_done == n_frames
    assert got.checksum_mismatches == 0 and got.masks_checked > 0
    for name in ("bits_left", "bits_right", "taxonomy", "iou_left", "iou_right"):
        a, b = getattr(got, name), getattr(ref, name)
        assert len(a) == len(b) == n_frames
        for i in range(n_frames):
            assert a[i].dtype == b[i].dtype and torch.equal(a[i], b[i]), (name, i)
    # the streaming sink saw every frame exactly once
    assert sorted(f for f, _, _ in blob["seen"]) == list(range(n_frames))
    for f, s, n in blob["seen"]:
        assert n == _n_prompts(f) and s == int(ref.masks(f, "left").sum())
    # the per-round batch sink saw every (round, source) block once, with consistent counts
    assert sum(len(c) for _, _, c, _, _ in blob["batches"]) == n_frames
    for src, lo, cnt, pb, ph in blob["batches"]:
        assert cnt == [_n_prompts(f) for f in range(lo, lo + len(cnt))] and pb == ph == sum(cnt)
        assert clip.shard_range(n_frames, src, world)[0] <= lo < clip.shard_range(n_frames, src, world)[1]


def test_zero_frames_and_bad_arguments():
    assert _single(0, 4).frames_done == 0
    with pytest.raises(ValueError):
        clip.segment_clip(_fake_segment_bits, _load, 4, (4, 4), HW, micro_batch=0, device=torch.device("cpu"))
    with pytest.raises(ValueError, match="max_prompts_per_frame"):
        clip.segment_clip(_fake_segment_bits, _load, 4, (4, 4), HW, micro_batch=2, max_prompts_per_frame=1,
                          device=torch.device("cpu"))


def test_corrupted_transfer_is_detected():
    """The checksum verification is live: a producer-side checksum that does not match the bytes counts as a mismatch."""
    calls = {"n": 0}

    def lying_checksum(data, out=None):
        calls["n"] += 1
        v = _checksum(data, out)
        if calls["n"] == 1:          # first producer-side call: claim a different checksum for mask 0
            v[0] += 1
        return v

    res = clip.segment_clip(_fake_segment_bits, _load, 2, (4, 4), HW, micro_batch=2, max_prompts_per_frame=3,
                            device=torch.device("cpu"), checksum_fn=lying_checksum)
    assert res.checksum_mismatches == 1


def test_consumer_errors_surface():
    def bad_sink(*a):
        raise RuntimeError("sink failed")

    with pytest.raises(RuntimeError, match="sink failed"):
        clip.segment_clip(_fake_segment_bits, _load, 3, (4, 4), HW, micro_batch=2, max_prompts_per_frame=3,
                          device=torch.device("cpu"), sink=bad_sink)


def test_slot_layout_is_aligned_and_round_trips():
    lay = clip.SlotLayout(8, 3, (1080, 1920))
    assert lay.Wb == 240 and lay.nbytes % 16 == 0 and lay.o_heads % 16 == 0 and lay.o_sums % 8 == 0
    assert lay.nbytes < 12.5e6          # 8 frames x 3 prompts x 2 sides at 1080p: 12.4 MB (99.5 MB as u8)
    buf = torch.zeros((lay.nbytes,), dtype=torch.uint8)
    bits, heads, counts, sums = lay.views(buf)
    bits[1, 5, 7, 3] = 0x81
    heads[2, 23, 3] = 1.5
    counts[7] = 3
    sums[47] = -1
    b2, h2, c2, s2 = lay.views(buf.clone())
    assert b2[1, 5, 7, 3] == 0x81 and h2[2, 23, 3] == 1.5 and c2[7] == 3 and s2[47] == -1
    m = torch.zeros((2, 5, 13), dtype=torch.uint8)
    m[1, 4, 12] = m[0, 0, 0] = m[0, 2, 8] = 1
    assert torch.equal(clip.unpack_bits(torch.from_numpy(A.pack_bits(m.numpy())), 13), m)
