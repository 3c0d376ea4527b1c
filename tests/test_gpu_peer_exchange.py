"""The peer-memory exchange (ob2_xchg_*) with `world` ranks emulated on one
GPU: every rank's inbox lives in the same device, so "peer stores" are plain
stores and the phases are ordered by the stream.  Must equal one set fed the
ranks' batches in rank order -- the same oracle as the NCCL path."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('dedup_grid', [0, 5])     # 5: the inbox dedup as a walking grid of five CTAs over the lanes
@pytest.mark.parametrize('world', [1, 2, 8])
def test_peer_exchange_emulated_on_one_gpu(world, dedup_grid):
    from orion_b200 import spaces
    from orion_b200.dist import PeerExchange, bucket_capacity
    from orion_b200.history import TrialHistory
    space = spaces.rosen2()
    n = 30000
    rng = np.random.RandomState(100 + world)
    history_fp = rng.randint(1, 2 ** 62, size=(n, 2), dtype=np.int64)
    pool = rng.randint(1, 2 ** 62, size=(world * n // 2, 2), dtype=np.int64)
    batches = []
    for _ in range(world):
        b = pool[rng.randint(0, len(pool), size=n)].copy()
        hit = rng.random_sample(n) < 0.15
        b[hit] = history_fp[rng.randint(0, n, size=int(hit.sum()))]
        b[rng.random_sample(n) < 0.02] = 0
        batches.append(b)
    cap = bucket_capacity(n, world)
    ranks = [PeerExchange(r, world, cap, 'cuda') for r in range(world)]
    regions = [x.region() for x in ranks]
    for x in ranks:
        if world > 1:
            x.set_peers(regions)
    shards = [TrialHistory(space, capacity=8 * n) for _ in range(world)]
    for shard in shards:
        shard.set_dedup_grid(dedup_grid)

    # history load through the exchange (insert_only)
    parts = np.array_split(history_fp, world)
    for r, x in enumerate(ranks):
        x.send(torch.from_numpy(np.ascontiguousarray(parts[r])).cuda(), len(parts[r]), 0)
    for r, x in enumerate(ranks):
        x.dedup(shards[r], insert_only=True)
    assert sum(len(s) for s in shards) == n

    for round_ in range(2):                       # second round: everything is known by then
        dev = [torch.from_numpy(b).cuda() for b in batches]
        for r, x in enumerate(ranks):
            x.send(dev[r], n, r * n)
        for r, x in enumerate(ranks):
            assert max(x.send_counts()) <= cap
            x.dedup(shards[r])
        seen = set(map(tuple, history_fp.tolist())) if round_ == 0 else seen
        for r, x in enumerate(ranks):
            keep = torch.full((n,), 7, dtype=torch.uint8, device='cuda')
            x.collect(n, keep)
            want = np.zeros(n, dtype=np.uint8)
            for i, (lo, hi) in enumerate(batches[r]):
                if lo != 0 and (int(lo), int(hi)) not in seen:
                    seen.add((int(lo), int(hi)))
                    want[i] = 1
            assert np.array_equal(keep.cpu().numpy(), want), 'round %d rank %d' % (round_, r)
        if round_ == 1:
            assert not any(w for w in want)
    assert sum(len(s) for s in shards) == len(seen)


def test_peer_exchange_lane_overflow_is_visible():
    from orion_b200.dist import PeerExchange
    x = PeerExchange(0, 1, 8, 'cuda')
    fp = torch.from_numpy(np.random.RandomState(0).randint(1, 2 ** 62, size=(100, 2), dtype=np.int64)).cuda()
    x.send(fp, 100, 0)
    assert x.send_counts() == [100]              # > cap = 8: the caller must treat the batch as overflowed
