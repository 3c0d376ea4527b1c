"""The multi-GPU dedup exchange (orion_b200/dist.py) with world_size 2 on CPU.

The choreography under test is the product's: bucket by owner -> all_to_all of
(fingerprint, global order key) -> owner-side first-wins dedup -> all_to_all of
verdicts -> scatter.  Only the three kernels are replaced by numpy stand-ins
(there is no GPU here); the GPU tests check those kernels against the same
semantics.  Expected result: exactly what ONE set would answer for the two
ranks' batches concatenated in rank order.
"""
import os
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from orion_b200.dist import ShardedHistory, bucket_capacity


class NumpyOps(object):
    """CPU stand-ins with the semantics of ob2_bucket_by_owner / ob2_set_dedup /
    ob2_set_insert / ob2_scatter_verdict."""

    def bucket(self, fp, n, key_base, world, cap, out):
        out_fp, out_key, out_src, counts = out
        f = fp.numpy()[:n]
        owner = f[:, 1].view(np.uint64) % np.uint64(world)
        out_fp.zero_()
        for r in range(world):
            rows = np.nonzero((owner == r) & (f[:, 0] != 0))[0]
            counts[r] = len(rows)
            rows = rows[:cap]
            out_fp[r, :len(rows)] = torch.from_numpy(f[rows])
            out_key[r, :len(rows)] = torch.from_numpy((key_base + rows).astype(np.int32))
            out_src[r, :len(rows)] = torch.from_numpy(rows.astype(np.int32))

    def dedup(self, history, fp, order, keep):
        f = fp.numpy()
        keys = order.numpy().view(np.uint32) if order is not None else np.arange(len(f), dtype=np.uint32)
        keep.zero_()
        for i in np.argsort(keys, kind='stable'):
            if f[i, 0] == 0:
                continue
            k = (int(f[i, 0]), int(f[i, 1]))
            if k not in history.rows:
                history.rows.add(k)
                keep[i] = 1

    def insert(self, history, fp):
        for lo, hi in fp.numpy():
            if lo != 0:
                history.rows.add((int(lo), int(hi)))

    def scatter(self, out_src, verdict, out_fp, keep):
        valid = out_fp.numpy()[:, 0] != 0
        keep.numpy()[out_src.numpy()[valid]] = verdict.numpy()[valid]


class FakeShard(object):
    device = torch.device('cpu')

    def __init__(self):
        self.rows = set()


def _make_batches(world, n, seed=0, distinct=None):
    """Per-rank fingerprints with duplicates inside a rank, across ranks and
    against the history.  ``distinct`` caps the number of different rows (a
    small discrete space): equal rows share an owner, so lanes overflow their
    nominal capacity."""
    rng = np.random.RandomState(seed)
    history = rng.randint(1, 2 ** 62, size=(n, 2), dtype=np.int64)
    pool = rng.randint(1, 2 ** 62, size=(distinct or world * n // 2, 2), dtype=np.int64)      # forces repeats
    batches = []
    for _ in range(world):
        b = pool[rng.randint(0, len(pool), size=n)].copy()
        hit = rng.random_sample(n) < 0.2
        b[hit] = history[rng.randint(0, n, size=int(hit.sum()))]
        batches.append(b)
    return history, batches


def _expected(history, batches):
    seen = set((int(a), int(b)) for a, b in history)
    out = []
    for b in batches:
        keep = np.zeros(len(b), dtype=np.uint8)
        for i, (lo, hi) in enumerate(b):
            k = (int(lo), int(hi))
            if k not in seen:
                seen.add(k)
                keep[i] = 1
        out.append(keep)
    return out


def _worker(rank, world, init_file, n, result_dir, distinct=None):
    dist.init_process_group('gloo', init_method='file://' + init_file, rank=rank, world_size=world)
    try:
        history, batches = _make_batches(world, n, distinct=distinct)
        sharded = ShardedHistory(FakeShard(), ops=NumpyOps(), device='cpu')
        assert sharded.world == world and sharded.rank == rank
        # history load: every rank contributes a slice, rows land on their owners
        part = np.array_split(history, world)[rank]
        pad = np.zeros((n, 2), dtype=np.int64)
        pad[:len(part)] = part
        sharded.insert(torch.from_numpy(pad), n)
        owned = [k for k in sharded.history.rows]
        assert all(np.uint64(hi) % np.uint64(world) == rank for _lo, hi in owned)
        # two batches in a row (the second sees what the first kept)
        fp = torch.from_numpy(batches[rank])
        keep1 = sharded.dedup(fp, n).numpy().copy()
        keep2 = sharded.dedup(fp, n).numpy().copy()
        np.save(os.path.join(result_dir, 'keep1_%d.npy' % rank), keep1)
        np.save(os.path.join(result_dir, 'keep2_%d.npy' % rank), keep2)
        np.save(os.path.join(result_dir, 'owned_%d.npy' % rank), np.array(len(owned)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_sharded_dedup_world2_matches_single_set():
    world, n = 2, 600
    with tempfile.TemporaryDirectory() as tmp:
        init_file = os.path.join(tmp, 'rendezvous')
        mp.spawn(_worker, args=(world, init_file, n, tmp), nprocs=world, join=True)
        history, batches = _make_batches(world, n)
        want = _expected(history, batches)
        owned = 0
        for r in range(world):
            keep1 = np.load(os.path.join(tmp, 'keep1_%d.npy' % r))
            keep2 = np.load(os.path.join(tmp, 'keep2_%d.npy' % r))
            assert np.array_equal(keep1[:n], want[r]), 'rank %d' % r
            assert not keep2[:n].any()          # everything is history by now
            owned += int(np.load(os.path.join(tmp, 'owned_%d.npy' % r)))
        assert owned == len(set(map(tuple, history.tolist())))
        assert 0 < sum(int(w.sum()) for w in want) < world * n


@pytest.mark.timeout(180)
def test_sharded_dedup_world2_skewed_owners_fall_back_to_worst_case_lanes():
    """Three distinct rows: one owner receives far more than mean + slack."""
    world, n, distinct = 2, 400, 3
    with tempfile.TemporaryDirectory() as tmp:
        init_file = os.path.join(tmp, 'rendezvous')
        mp.spawn(_worker, args=(world, init_file, n, tmp, distinct), nprocs=world, join=True)
        history, batches = _make_batches(world, n, distinct=distinct)
        want = _expected(history, batches)
        for r in range(world):
            keep1 = np.load(os.path.join(tmp, 'keep1_%d.npy' % r))
            assert np.array_equal(keep1[:n], want[r]), 'rank %d' % r
        assert sum(int(w.sum()) for w in want) <= distinct


def test_single_rank_path_skips_the_exchange():
    history, batches = _make_batches(1, 300, seed=3)
    sharded = ShardedHistory(FakeShard(), ops=NumpyOps(), device='cpu')
    assert sharded.world == 1
    sharded.insert(torch.from_numpy(history), len(history))
    keep = sharded.dedup(torch.from_numpy(batches[0]), 300)
    assert np.array_equal(keep.numpy()[:300], _expected(history, batches)[0])


def test_bucket_capacity_has_slack():
    for n, world in [(100, 2), (10 ** 6, 8), (12_500_000, 8)]:
        cap = bucket_capacity(n, world)
        assert cap > n / world * 1.05 and cap < n / world * 1.2 + 200
