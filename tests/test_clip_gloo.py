"""world_size-2 (and 3) gloo tests of the frame-sharded clip path (2handedafforder_b200/clip.py).

The exchange logic (shard ranges, double-buffered asynchronous gather, ragged shards, variable prompt counts, the
streaming sink) is host code and is exercised here on CPU with a deterministic stand-in for `B200Sam.segment`; the
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


def _fake_segment(frames, segs, resize_list, original_list, threshold=0.0):
    """Deterministic stand-in with the output contract of B200Sam.segment."""
    H, W = original_list[0]
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
    return {"masks_left_packed": st(ml, (H, W)).to(torch.uint8), "masks_right_packed": st(mr, (H, W)).to(torch.uint8),
            "taxonomy": st(tx, (4,)), "iou_left": st(il, (4,)), "iou_right": st(ir, (4,))}


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
        seen = []
        res = clip.segment_clip(_fake_segment, _load, n_frames, (4, 4), HW, micro_batch=mb, max_prompts_per_frame=3,
                                device=torch.device("cpu"),
                                sink=lambda f, *a: seen.append((f, int(a[0].sum()), a[0].shape[0])))
        if rank == 0:
            torch.save({"res": res, "seen": seen}, path)
        else:
            assert res.frames_done == 0 and not res.masks_left
    finally:
        dist.destroy_process_group()


def _single(n_frames, mb):
    return clip.segment_clip(_fake_segment, _load, n_frames, (4, 4), HW, micro_batch=mb, max_prompts_per_frame=3,
                             device=torch.device("cpu"))


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
    for i in range(5):
        direct = _fake_segment(_frame(i)[None], [_seg(i)], [(4, 4)], [HW])
        assert torch.equal(res.masks_left[i], direct["masks_left_packed"])
        assert torch.equal(res.iou_right[i], direct["iou_right"])
        assert res.masks_left[i].shape == (_n_prompts(i), *HW)


@pytest.mark.parametrize("world,n_frames,mb", [(2, 13, 4), (2, 8, 8), (3, 7, 2), (2, 1, 4)])
def test_gathered_clip_is_identical_to_one_rank(world, n_frames, mb, tmp_path):
    path = str(tmp_path / "rank0.pt")
    mp.spawn(_worker, args=(world, _free_port(), n_frames, mb, path), nprocs=world, join=True)
    blob = torch.load(path, weights_only=False)
    got, ref = blob["res"], _single(n_frames, mb)
    assert got.frames_done == n_frames
    for name in ("masks_left", "masks_right", "taxonomy", "iou_left", "iou_right"):
        a, b = getattr(got, name), getattr(ref, name)
        assert len(a) == len(b) == n_frames
        for i in range(n_frames):
            assert a[i].dtype == b[i].dtype and torch.equal(a[i], b[i]), (name, i)
    # the streaming sink saw every frame exactly once
    assert sorted(f for f, _, _ in blob["seen"]) == list(range(n_frames))
    for f, s, n in blob["seen"]:
        assert n == _n_prompts(f) and s == int(ref.masks_left[f].sum())


def test_zero_frames_and_bad_arguments():
    assert _single(0, 4).frames_done == 0
    with pytest.raises(ValueError):
        clip.segment_clip(_fake_segment, _load, 4, (4, 4), HW, micro_batch=0, device=torch.device("cpu"))
    with pytest.raises(ValueError, match="max_prompts_per_frame"):
        clip.segment_clip(_fake_segment, _load, 4, (4, 4), HW, micro_batch=2, max_prompts_per_frame=1,
                          device=torch.device("cpu"))
