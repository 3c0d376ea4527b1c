"""Edge cases: empty and degenerate inputs, special values, hash-slot
collisions, limits."""
import numpy as np
import pytest
import torch

from oracle import space_oracle as so
from tests import golden_util as gu

pytestmark = pytest.mark.gpu


def test_empty_space_and_zero_samples():
    from orion_b200.space import Real, Space
    space = Space()
    assert space.sample(3) == []
    assert (1, 2) not in space
    with pytest.raises(ValueError):
        space.compiled()
    space.register(Real('x', 'uniform', 0, 1))
    assert space.sample(0) == []
    b = space.sample_columns(0)
    assert b.n == 0 and b.to_numpy()[0].shape == (0,)
    assert space.contains_points([]) == []


def test_contains_special_values_match_numpy_semantics():
    """NaN is never inside; -inf is inside an unbounded-below Real (SURVEY hard
    part 5: `-inf in Real('norm')` is True in the reference); +inf never is."""
    from orion_b200.columns import ColumnBatch
    dims = [dict(name='g', kind='real', prior='norm', args=[0, 1]),
            dict(name='u', kind='real', prior='uniform', args=[-1, 2]),
            dict(name='i', kind='integer', prior='norm', args=[0, 5])]
    specs = gu.specs_from_json(dims)
    space = gu.product_space(dims)
    compiled = space.compiled()
    vals = [np.nan, -np.inf, np.inf, 0.0, -0.0, 1.0 - 1e-16, 1.0, -1.0, np.nextafter(-1.0, -2.0), 3.0, 2.5]
    n = len(vals)
    cols = [np.zeros(n), np.zeros(n), np.zeros(n)]
    got = {}
    for s in range(3):
        probe = [c.copy() for c in cols]
        probe[s] = np.asarray(vals, dtype=np.float64)
        batch = ColumnBatch.from_numpy(compiled, probe, int_as_f64=True)
        got[s] = compiled.contains(batch, int_as_f64=True).cpu().numpy().astype(bool)
        with np.errstate(invalid='ignore'):
            want = so.contains(specs, probe)
        assert np.array_equal(got[s], want), (s, got[s], want)
    assert list(got[0][:3]) == [False, True, False]          # norm: nan, -inf, +inf
    assert list(got[2][:3]) == [False, False, False]         # Integer: inf % 1 is nan -> not integral
    assert got[1][5] and not got[1][6] and got[1][7] and not got[1][8]   # half-open [-1, 1)


def test_u_equal_zero_and_almost_one():
    """u = 0 gives the lower edge (ndtri(0) = -inf is accepted by an unbounded
    normal); u = 1 - 2^-53 stays below the exclusive upper bound."""
    dims = [dict(name='g', kind='real', prior='norm', args=[0, 1]),
            dict(name='u', kind='real', prior='uniform', args=[-1, 2]),
            dict(name='l', kind='real', prior='reciprocal', args=[1e-5, 1]),
            dict(name='i', kind='integer', prior='uniform', args=[-100, 200]),
            dict(name='c', kind='categorical', categories=list(range(8)))]
    specs = gu.specs_from_json(dims)
    compiled = gu.product_space(dims).compiled()
    u = np.array([0.0, 1.0 - 2.0 ** -53, 0.5, 2.0 ** -53] + [0.25] * 28)
    U = torch.from_numpy(np.tile(u, (5, 1))).cuda()
    batch = compiled.sample_from_uniform(U)
    got = batch.to_numpy()
    want, ok = so.sample_from_uniform(specs, np.tile(u, (5, 1)))
    assert got[0][0] == -np.inf and np.array_equal(got[1], want[1])
    np.testing.assert_allclose(got[0][1:], want[0][1:], rtol=1e-6)
    np.testing.assert_allclose(got[2], want[2], rtol=1e-6)
    assert np.array_equal(got[3], want[3]) and np.array_equal(got[4], want[4])
    assert got[3][0] == -100 and got[3][1] == 99 and got[4][0] == 0 and got[4][1] == 7
    assert np.array_equal(batch.mask.cpu().numpy().astype(bool), ok)


def test_hash_set_slot_collisions_and_shared_low_words():
    """Keys that land on the same slot chain (equal low bits) and keys that
    share the whole low word but differ in the high word are distinct rows."""
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    hist = TrialHistory(spaces.rosen2(), n_slots=64)
    # 20 keys with identical (lo & 63): one probe chain of length 20
    chain = np.array([[(k << 6) | 5, 1000 + k] for k in range(1, 21)], dtype=np.int64)
    # 10 keys with the same lo and different hi
    same_lo = np.array([[777, k] for k in range(1, 11)], dtype=np.int64)
    fp = np.concatenate([chain, same_lo, chain[:5], same_lo[:3]])
    keep = hist.dedup_fingerprints(torch.from_numpy(fp).cuda()).cpu().numpy()
    assert list(keep) == [1] * 30 + [0] * 8
    assert len(hist) == 30
    found = hist.probe_fingerprints(torch.from_numpy(np.array([[777, 4], [777, 99], [(3 << 6) | 5, 1003],
                                                                [(3 << 6) | 5, 7]], dtype=np.int64)).cuda())
    assert list(found.cpu().numpy()) == [1, 0, 1, 0]
    # wrap-around of the probe sequence at the end of the table
    wrap = np.array([[63 + 64 * k, k] for k in range(1, 6)], dtype=np.int64)
    keep = hist.dedup_fingerprints(torch.from_numpy(wrap).cuda()).cpu().numpy()
    assert list(keep) == [1] * 5 and len(hist) == 35


def test_dedup_all_duplicates_and_all_padding():
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    hist = TrialHistory(spaces.rosen2(), capacity=1 << 14)
    same = torch.tensor([[42, 7]] * 5000, dtype=torch.int64).cuda()
    keep = hist.dedup_fingerprints(same).cpu().numpy()
    assert keep[0] == 1 and not keep[1:].any() and len(hist) == 1          # first index wins, deterministically
    order = torch.arange(5000, 0, -1, dtype=torch.int32).cuda()            # reversed keys: the LAST position wins
    same2 = torch.tensor([[43, 7]] * 5000, dtype=torch.int64).cuda()
    keep = hist.dedup_fingerprints(same2, order=order).cpu().numpy()
    assert keep[-1] == 1 and keep.sum() == 1
    pad = torch.zeros((100, 2), dtype=torch.int64).cuda()
    assert not hist.dedup_fingerprints(pad).cpu().numpy().any() and len(hist) == 2


def test_determinism_across_runs():
    """Same inputs -> same keep mask, whatever the thread interleaving."""
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    rng = np.random.RandomState(5)
    pool = rng.randint(1, 2 ** 62, size=(3000, 2), dtype=np.int64)
    fp = torch.from_numpy(pool[rng.randint(0, 3000, size=200_000)]).cuda()
    masks = []
    for _ in range(3):
        hist = TrialHistory(spaces.rosen2(), capacity=1 << 19)     # sized for the incoming batch, not the distinct keys
        masks.append(hist.dedup_fingerprints(fp).clone())
    assert torch.equal(masks[0], masks[1]) and torch.equal(masks[0], masks[2])
    first = np.unique(fp.cpu().numpy()[:, 0], return_index=True)[1]
    want = np.zeros(200_000, dtype=np.uint8)
    want[first] = 1
    assert np.array_equal(masks[0].cpu().numpy(), want)


def test_wide_space_descriptor_limits():
    """A space whose descriptor does not fit the 227 KiB staging limit (all the shared memory of one block) is
    refused with a clear error; a wide one below it works."""
    from orion_b200 import _lib
    from orion_b200.space import Real, Space
    space = Space()
    space.register(Real('x', 'uniform', 0, 1, shape=(1500,)))
    batch = space.sample_columns(100, seed=1)                   # 1500 slots x 152 B (descriptor + draw plan + run) = 223 KiB: fits
    assert bool(batch.mask[:100].all().item())
    big = Space()
    big.register(Real('x', 'uniform', 0, 1, shape=(1700,)))
    with pytest.raises(_lib.Ob2Error) as exc:
        big.compiled()
    assert '227 KiB' in str(exc.value)


def test_string_offered_to_a_numeric_dimension_raises_like_the_reference():
    """The reference compares with `point_ < high` (space.py:164): a str there is a
    TypeError, not the number it spells; a bool is a number for numpy."""
    from orion_b200.space import Integer, Real, Space
    dim = Real('x', 'uniform', 0, 10)
    with pytest.raises(TypeError):
        '3' in dim                                        # noqa: B015
    space = Space()
    space.register(dim)
    space.register(Integer('k', 'uniform', 0, 5))
    with pytest.raises(TypeError):
        ('3', 1) in space                                 # noqa: B015
    assert (True, 1) in space
    assert (3.0, 1) in space and (3.0, 1.5) not in space
