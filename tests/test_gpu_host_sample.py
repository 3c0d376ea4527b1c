"""ob2_sample_host: the small-n Space.sample path (kernel + copy + sync in one
library call) returns exactly what the tensor path returns."""
import numpy as np
import pytest

from tests import golden_util as gu

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('n', [0, 1, 10, 33, 4096])
def test_sample_host_equals_tensor_path(n):
    g = gu.load('bounded_reject')
    space = gu.product_space(g['dims'] + [dict(name='cc9', kind='categorical', categories=['a', 'b', 'c'], shape=[2])])
    compiled = space.compiled()
    cols, failed = compiled.sample_host(n, 77, base_index=5)
    batch = compiled.sample(max(n, 1), 77, base_index=5)
    for a, b in zip(cols, batch.to_numpy()):
        assert a.dtype == b.dtype and np.array_equal(a, b[:n])
    if n:
        assert np.array_equal(failed, batch.n_failed.cpu().numpy().astype(np.uint32))
    # returned columns are the caller's: a second call must not change them
    keep = [c.copy() for c in cols]
    compiled.sample_host(max(n, 1), 78)
    for a, b in zip(cols, keep):
        assert np.array_equal(a, b)


def test_sample_host_limits_and_errors():
    from orion_b200 import _lib, spaces
    from orion_b200.space import Real, Space
    compiled = spaces.rosen2().compiled()
    with pytest.raises(_lib.Ob2Error) as exc:
        compiled.sample_host(_lib.HOST_SAMPLE_MAX + 1, 1)
    assert 'ob2_sample' in str(exc.value)
    space = Space()
    space.register(Real('g', 'ncx2', 3.0, 1.0))            # no device inverse CDF: refuses, never falls back
    with pytest.raises(NotImplementedError):
        space.sample(1)
    improbable = gu.load('improbable')
    with pytest.raises(ValueError) as exc:
        gu.product_space(improbable['dims']).sample(8, seed=3)          # small n: the host path raises the same
    assert str(exc.value) == improbable['error']


def test_space_sample_same_rows_on_both_paths(monkeypatch):
    """On the Philox stream Space.sample switches paths at HOST_SAMPLE_MAX; a seed gives the same
    leading rows either way.  (The reference's own stream is dimension-major: its rows depend on n.)"""
    from orion_b200 import _lib, mt_stream, spaces
    monkeypatch.setattr(mt_stream, 'ENABLED', False)
    space = spaces.mixed64()
    small = space.sample(50, seed=9)
    big = space.sample(_lib.HOST_SAMPLE_MAX + 50, seed=9)
    assert small == big[:50]
    dim = space['u0'] if 'u0' in space else list(space.values())[0]
    assert dim.sample(7, seed=3) == dim.sample(_lib.HOST_SAMPLE_MAX + 7, seed=3)[:7]


@pytest.mark.parametrize('n', [1, 10, 700])
def test_contains_host_equals_tensor_path(n):
    from orion_b200.columns import ColumnBatch
    g = gu.load('bounded_reject')
    space = gu.product_space(g['dims'] + [dict(name='cc9', kind='categorical', categories=['a', 'b', 'c'], shape=[2])])
    compiled = space.compiled()
    cols = [c.copy() for c in compiled.sample_host(n, 5)[0]]
    rng = np.random.RandomState(n)
    for j, (_u, dt) in enumerate(compiled.col_meta):            # knock some values out of the space
        bad = rng.random_sample(n) < 0.1
        if dt == 'f8':
            cols[j] = np.where(bad, np.nan if j % 2 else 1e300, cols[j])
        elif dt == 'i8':
            cols[j] = np.where(bad, 0.5, cols[j].astype(np.float64))        # not integral
        else:
            cols[j] = np.where(bad, 99, cols[j]).astype(np.int32)
    got = compiled.contains_host(cols, n, int_as_f64=True)
    batch = ColumnBatch.from_numpy(compiled, cols, int_as_f64=True)
    want = compiled.contains(batch, int_as_f64=True).cpu().numpy().astype(bool)
    assert got.dtype == bool and np.array_equal(got, want) and (not want.all() or n == 1)


def test_points_in_space_small_and_bulk_agree():
    from orion_b200 import _lib, spaces
    space = spaces.mixed64()
    pts = space.sample(40, seed=1)
    bad = list(pts[3])
    bad[0] = 7.0                                            # uniform(-1, 1)
    pts[3] = tuple(bad)
    small = space.contains_points(pts)
    assert small == [i != 3 for i in range(40)]
    many = pts * ((_lib.HOST_SAMPLE_MAX // 40) + 2)         # > HOST_SAMPLE_MAX: the device-batch path
    assert space.contains_points(many) == small * ((_lib.HOST_SAMPLE_MAX // 40) + 2)
    assert pts[0] in space and pts[3] not in space
    assert pts[0][0] in space['u0'] if 'u0' in space else True


def test_sample_dedup_host_equals_tensor_path():
    """ob2_sample_dedup_host == sample + fingerprint + TrialHistory.dedup."""
    from orion_b200.history import TrialHistory
    from orion_b200.space import Categorical, Integer, Space
    space = Space()                                  # small discrete space: duplicates are certain
    space.register(Integer('i', 'uniform', 0, 6))
    space.register(Categorical('c', ['a', 'b', 'c']))
    compiled = space.compiled()
    h_host, h_dev = TrialHistory(space, capacity=4096), TrialHistory(space, capacity=4096)
    for call, n in enumerate([1, 7, 64, 500]):
        cols, failed, keep = compiled.sample_host(n, 100 + call, history=h_host)
        batch = compiled.sample(n, 100 + call, with_fingerprint=True)
        want_keep = h_dev.dedup(batch).cpu().numpy().astype(bool)
        assert keep.dtype == bool and np.array_equal(keep, want_keep)
        for a, b in zip(cols, batch.to_numpy()):
            assert np.array_equal(a, b)
        assert not failed.any()
        assert len(h_host) == len(h_dev)
    assert len(h_host) == 18                         # every (i, c) pair seen, each kept exactly once


def test_random_suggest_with_dedup_small_calls():
    from orion_b200.algo.random import Random
    from orion_b200.space import Categorical, Integer, Space
    space = Space()
    space.register(Integer('i', 'uniform', 0, 6))
    space.register(Categorical('c', ['a', 'b', 'c']))
    algo = Random(space, seed=5, dedup=True)
    seen = []
    for _ in range(6):
        pts = algo.suggest(3)
        assert len(pts) == 3 and all(p in space for p in pts)
        seen.extend((int(p[0]), p[1]) for p in pts)
    assert len(set(seen)) == 18                      # never the same point twice: the space is exhausted exactly
    assert algo.suggest(1) == []                     # nothing new left (bounded top-up gives up)
    auto = Random(space, seed=5, dedup=True, history_capacity=8)
    auto.history.auto_grow = True
    assert len(auto.suggest(12)) == 12               # the tiny table grows instead of failing
