"""ob2_set_set_dedup_grid: the dedup's first pass as a small walking grid (the form that shares the SMs with a
kernel on another stream) must give the verdicts of the default one-thread-per-entry grid, bit for bit -- the
first-wins rule is decided by order keys, not by which thread got there first."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _batch(rng, n, history):
    fp = rng.randint(1, 2 ** 62, size=(n, 2), dtype=np.int64)
    dup = rng.random_sample(n) < 0.05                       # duplicates of earlier rows of the same batch
    src = (rng.random_sample(n) * np.arange(n)).astype(np.int64)
    fp[dup] = fp[src[dup]]
    hit = rng.random_sample(n) < 0.10                       # rows the history already holds
    fp[hit] = history[rng.randint(0, len(history), size=int(hit.sum()))]
    fp[rng.random_sample(n) < 0.01] = 0                     # padding entries
    return fp


@pytest.mark.parametrize('grid', [1, 7, 222, 60000])
def test_walking_grid_gives_the_default_grid_s_verdicts(grid):
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    space = spaces.rosen2()
    rng = np.random.RandomState(grid)
    n = 300_000
    history = rng.randint(1, 2 ** 62, size=(n, 2), dtype=np.int64)
    tables = [TrialHistory(space, capacity=8 * n) for _ in range(2)]
    tables[1].set_dedup_grid(grid)
    for t in tables:
        t.insert_fingerprints(torch.from_numpy(history).cuda(), n)
    for _round in range(3):
        fp = torch.from_numpy(_batch(rng, n, history)).cuda()
        keeps = [t.dedup_fingerprints(fp, n).clone() for t in tables]
        assert torch.equal(keeps[0], keeps[1])
        assert 0 < int(keeps[0][:n].sum().item()) < n
        assert len(tables[0]) == len(tables[1])


def test_bad_grid_is_refused():
    from orion_b200 import _lib, spaces
    from orion_b200.history import TrialHistory
    t = TrialHistory(spaces.rosen2(), capacity=1024)
    with pytest.raises(_lib.Ob2Error):
        t.set_dedup_grid(-1)
    with pytest.raises(_lib.Ob2Error):
        t.set_dedup_grid(1 << 20)


def test_grid_setting_survives_growth():
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    rng = np.random.RandomState(3)
    n = 50_000
    fp = torch.from_numpy(rng.randint(1, 2 ** 62, size=(n, 2), dtype=np.int64)).cuda()
    small = TrialHistory(spaces.rosen2(), capacity=1 << 12, auto_grow=True)     # far too small: grows on the way
    small.set_dedup_grid(3)
    keep = small.dedup_fingerprints(fp, n)
    assert int(keep[:n].sum().item()) == n == len(small)
    assert small._dedup_grid == 3
    again = small.dedup_fingerprints(fp, n)
    assert int(again[:n].sum().item()) == 0
