"""History table growth (ob2_set_rehash) and snapshot / restore (ob2_set_export)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _fps(rng, n):
    return rng.randint(1, 2 ** 62, size=(n, 2), dtype=np.int64)


def test_grow_keeps_every_fingerprint_and_dedup_semantics():
    from orion_b200 import _lib, spaces
    from orion_b200.history import TrialHistory
    rng = np.random.RandomState(0)
    hist = TrialHistory(spaces.rosen2(), n_slots=1 << 12)
    first = _fps(rng, 3000)
    hist.insert_fingerprints(torch.from_numpy(first).cuda())
    second = _fps(rng, 500)
    keep = hist.dedup_fingerprints(torch.from_numpy(second).cuda()).cpu().numpy()
    assert keep.all() and len(hist) == 3500
    with pytest.raises(_lib.Ob2Error):                         # 3500 + 1000 > 7/8 * 4096
        hist.insert_fingerprints(torch.from_numpy(_fps(rng, 1000)).cuda())
    hist.grow()
    assert hist.n_slots == 1 << 13 and len(hist) == 3500
    everything = torch.from_numpy(np.concatenate([first, second])).cuda()
    assert bool(hist.probe_fingerprints(everything).all().item())
    # after growth all earlier keys are history: duplicates of them are dropped, new keys kept, first wins
    third = _fps(rng, 2000)
    third[::4] = first[:500]
    third[1::4] = third[2::4]                                  # in-batch duplicates (the earlier position wins)
    got = hist.dedup_fingerprints(torch.from_numpy(third).cuda()).cpu().numpy()
    want = np.ones(2000, dtype=np.uint8)
    want[::4] = 0
    want[2::4] = 0
    assert np.array_equal(got, want)
    with pytest.raises(ValueError):
        hist.grow(1 << 13)


def test_auto_grow_under_dedup_load():
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    rng = np.random.RandomState(1)
    hist = TrialHistory(spaces.rosen2(), n_slots=1 << 10, auto_grow=True)
    seen = set()
    for step in range(12):
        batch = _fps(rng, 700)
        if step:
            batch[:100] = prev[:100]                           # noqa: F821  (copies of the previous batch)
        got = hist.dedup_fingerprints(torch.from_numpy(batch).cuda()).cpu().numpy()
        want = np.zeros(700, dtype=np.uint8)
        for i, (lo, hi) in enumerate(batch):
            if (int(lo), int(hi)) not in seen:
                seen.add((int(lo), int(hi)))
                want[i] = 1
        assert np.array_equal(got, want), step
        prev = batch
    assert len(hist) == len(seen) and hist.n_slots > 1 << 10


def test_snapshot_and_restore_into_a_different_table_size():
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    rng = np.random.RandomState(2)
    space = spaces.mixed64()
    hist = TrialHistory(space, capacity=1 << 14)
    batch = space.sample_columns(5000, seed=1, with_fingerprint=True)
    hist.dedup(batch)
    state = hist.state_dict()
    assert state['fingerprints'].shape == (5000, 2) and not state['fingerprints'].is_cuda
    lo, hi = batch.fingerprints_numpy()
    stored = set(map(tuple, state['fingerprints'].numpy().view(np.uint64).tolist()))
    assert stored == set(zip(lo.tolist(), hi.tolist()))
    restored = TrialHistory(space, capacity=1 << 16)
    restored.load_state_dict(state)
    assert len(restored) == 5000
    again = space.sample_columns(5000, seed=1, with_fingerprint=True)
    assert not bool(restored.dedup(again).any().item())
    empty = TrialHistory(space, capacity=64).state_dict()
    assert empty['fingerprints'].shape == (0, 2)
