"""ob2_suggest (sample -> transform -> reverse -> fingerprint in one kernel)
must be bit-identical to the staged kernels it fuses, and to the oracle."""
import numpy as np
import pytest
import torch

from oracle import space_oracle as so
from tests import golden_util as gu

pytestmark = pytest.mark.gpu


def _staged(compiled, n, seed, index=None):
    tb = compiled.empty_columns(n, transformed=True)
    batch = compiled.sample(n, seed, index=index, transformed_out=tb)
    back = compiled.reverse(tb, with_fingerprint=True)
    return batch, tb, back


def _same(a, b, n):
    for x, y in zip(a.to_numpy(), b.to_numpy()):       # (buffers differ in the stride padding)
        assert x.dtype == y.dtype and np.array_equal(x, y)


@pytest.mark.parametrize('requirement', ['real', 'integer', 'numerical', None])
@pytest.mark.parametrize('quantize', ['round', 'floor'])
@pytest.mark.parametrize('n', [1, 2, 3001])
def test_suggest_equals_staged_kernels(requirement, quantize, n):
    """Includes rows that went through the redraw path (bounded_reject) and the
    lossy Quantize round trip of requirement 'integer'."""
    from orion_b200.transformer import build_required_space
    g = gu.load('bounded_reject')
    extra = [dict(d, name='m_' + d['name']) for d in gu.load('mixed_scalar')['dims'][5:]]
    dims = g['dims'] + extra + [dict(name='b2', kind='categorical', categories=['x', 'y']),
                                dict(name='b1', kind='categorical', categories=['only']),
                                dict(name='sh', kind='categorical', categories=[1, 2, 3], shape=[2])]
    space = gu.product_space(dims)
    compiled = build_required_space(requirement, space, quantize=quantize).compiled()
    seed = 123
    b0, t0, k0 = _staged(compiled, n, seed)
    b1, t1, k1 = compiled.suggest(n, seed)
    _same(b1, b0, n)
    _same(t1, t0, n)
    _same(k1, k0, n)
    assert torch.equal(k1.fp[:n], k0.fp[:n])
    assert torch.equal(b1.mask[:n], b0.mask[:n])
    assert torch.equal(b1.n_failed, b0.n_failed)
    if n % 2:
        assert int(k1.fp[n, 0].item()) == 0            # padding entry of the odd tail
    # ... and the fingerprints are the oracle's over the handed-out rows
    lo, hi = k1.fingerprints_numpy()
    wlo, whi = so.fingerprint(k1.to_numpy())
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)


def test_suggest_matches_oracle_mixed64():
    from orion_b200 import spaces
    from orion_b200.transformer import build_required_space
    from tests.spec_util import specs_from_space
    space = spaces.mixed64()
    specs = specs_from_space(space)
    compiled = build_required_space('real', space).compiled()
    n, seed, base = 20_001, 20240003, (1 << 33) + 5
    batch, tb, back = compiled.suggest(n, seed, base_index=base)
    want, failed = so.sample_philox(specs, n, seed, base_index=base)
    assert not failed.any()
    # explicit counters replay the same rows (and agree with the staged kernels)
    index = torch.arange(base, base + n, dtype=torch.int64).cuda().flip(0).contiguous()
    rb, rt, rk = compiled.suggest(n, seed, index=index)
    sb, st, sk = _staged(compiled, n, seed, index=index)
    _same(rb, sb, n)
    _same(rk, sk, n)
    assert torch.equal(rk.fp[:n], sk.fp[:n]) and torch.equal(rk.fp[:n], back.fp[:n].flip(0))
    for s, (a, b) in enumerate(zip(batch.to_numpy(), want)):
        if specs_kind(specs, s) == 'real':
            assert np.allclose(a, b, rtol=1e-6, atol=0), s   # loguniform / normal: <= 1e-6 relative
        else:
            assert np.array_equal(a, b), s
    got = batch.to_numpy()
    twant = so.transform(specs, 'real', got)
    for a, b in zip(tb.to_numpy(), twant):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    for a, b in zip(back.to_numpy(), so.reverse(specs, 'real', twant)):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    assert bool(batch.mask[:n].all())


def specs_kind(specs, slot):
    s = 0
    for spec in specs:
        width = int(np.prod(spec.shape)) if spec.shape else 1
        if slot < s + width:
            return spec.kind
        s += width
    raise IndexError(slot)


def test_suggest_extended_priors_and_improbable_counts():
    from orion_b200.space import Integer, Real, Space
    from orion_b200.transformer import build_required_space
    space = Space()
    space.register(Real('e', 'expon', 0.0, 2.0))
    space.register(Real('n', 'norm', 0.0, 1.0, low=2.5, high=2.6))       # misses 4 times now and then
    space.register(Integer('i', 'uniform', -3, 9))
    compiled = build_required_space('real', space).compiled()
    n, seed = 40_000, 11
    b0, t0, k0 = _staged(compiled, n, seed)
    b1, t1, k1 = compiled.suggest(n, seed)
    assert int(b0.n_failed.sum().item()) > 0
    assert torch.equal(b1.n_failed, b0.n_failed)
    _same(b1, b0, n)
    _same(t1, t0, n)
    _same(k1, k0, n)
    assert torch.equal(k1.fp[:n], k0.fp[:n]) and torch.equal(b1.mask[:n], b0.mask[:n])


def test_transformed_space_suggest_columns():
    from orion_b200 import spaces
    from orion_b200.transformer import build_required_space
    tspace = build_required_space('real', spaces.mixed64())
    n, seed = 1000, 42
    batch, tbatch, back = tspace.suggest_columns(n, seed=seed)
    b0, t0 = tspace.sample_columns(n, seed=seed)
    _same(batch, b0, n)
    _same(tbatch, t0, n)
    _same(back, tspace.reverse_columns(t0), n)
    assert bool(tspace.contains_columns(tbatch).all())
    rows = back.to_numpy()
    lo, hi = back.fingerprints_numpy()
    wlo, whi = so.fingerprint(rows)
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)


@pytest.mark.parametrize('requirement', ['real', 'integer', None])
def test_suggest_with_back_aliased_to_cols_keeps_the_handed_out_rows(requirement):
    """back == cols: one buffer, holding the rows as handed out (which differ
    from the sampled ones only through a lossy Quantize), same fingerprints,
    same transformed columns -- redraw rows included."""
    from orion_b200.transformer import build_required_space
    g = gu.load('bounded_reject')
    extra = [dict(d, name='m_' + d['name']) for d in gu.load('mixed_scalar')['dims'][5:]]
    space = gu.product_space(g['dims'] + extra + [dict(name='sh', kind='categorical', categories=[1, 2, 3], shape=[2])])
    compiled = build_required_space(requirement, space).compiled()
    for n in (1, 3001):
        b0, t0, k0 = compiled.suggest(n, 123)
        one = compiled.empty_columns(n)
        b1, t1, k1 = compiled.suggest(n, 123, out=one, back_out=one)
        assert b1 is k1
        _same(k1, k0, n)
        _same(t1, t0, n)
        assert torch.equal(k1.fp[:n], k0.fp[:n]) and torch.equal(b1.mask[:n], b0.mask[:n])
        assert torch.equal(b1.n_failed, b0.n_failed)


def test_pipeline_fused_agrees_with_staged():
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    from orion_b200.pipeline import CandidatePipeline
    from orion_b200.transformer import build_required_space
    compiled = build_required_space('real', spaces.mixed64()).compiled()
    n = 50_000
    rng = np.random.RandomState(3)
    index = [torch.from_numpy(rng.randint(0, 2 * n, size=n).astype(np.int64)).cuda() for _ in range(4)]
    kept, rows = {}, {}
    for mode in ('suggest', 'reverse'):
        pipe = CandidatePipeline(compiled, TrialHistory(compiled, capacity=8 * n), n, fingerprint_in=mode)
        out, back = [], []
        for k in range(4):
            slot = pipe.submit(5, index=index[k])
            pipe.drain()
            torch.cuda.synchronize()
            out.append(slot.batch.keep[:n].clone())
            back.append([torch.from_numpy(c) for c in slot.back.to_numpy()])
        kept[mode], rows[mode] = out, back
    for a, b in zip(kept['suggest'], kept['reverse']):
        assert torch.equal(a, b)
    for ra, rb in zip(rows['suggest'], rows['reverse']):
        for a, b in zip(ra, rb):
            assert torch.equal(a, b)


@pytest.mark.parametrize('requirement', ['real', 'integer'])
def test_random_over_a_transformed_space(requirement):
    """Random wrapped the way PrimaryAlgo is meant to wrap it: its space is the
    TransformedSpace, its points are transformed points, dedup is on the rows
    as handed out."""
    from orion_b200.algo.random import Random
    from orion_b200.space import Categorical, Integer, Real, Space
    from orion_b200.transformer import build_required_space
    space = Space()
    space.register(Integer('i', 'uniform', 0, 4))
    space.register(Categorical('c', ['a', 'b', 'c']))
    space.register(Real('x', 'uniform', 0, 2))
    # (floor: Quantize by rounding could hand out x = 2.0, outside the original [0, 2) -- Appendix B.2's
    # `reverse(p) in original` would then reject the point)
    tspace = build_required_space(requirement, space, quantize='floor')
    algo = Random(tspace, seed=3)
    pts = algo.suggest(5)
    assert len(pts) == 5 and all(p in tspace for p in pts)
    assert all(tspace.reverse(p) in space for p in pts)
    tb = algo.suggest_columns(1000)
    assert tb.transformed and bool(tb.mask[:1000].all()) and tb.original.n == 1000
    for a, b in zip(tspace.reverse_columns(tb).to_numpy(), tb.original.to_numpy()):
        assert np.array_equal(a, b)
    # dedup: 'integer' floors x to {0, 1}: 4 * 3 * 2 = 24 distinct handed-out rows, never one twice
    dd = Random(tspace, seed=4, dedup=True)
    if requirement == 'integer':
        seen = []
        for _ in range(8):
            got = dd.suggest(3)
            seen.extend(tuple(np.asarray(v).tolist() if hasattr(v, 'tolist') else v for v in tspace.reverse(p)) for p in got)
        assert len(seen) == 24 and len(set(seen)) == 24
        assert dd.suggest(1) == []
    else:
        got = dd.suggest(50)
        assert len(got) == 50
        dd.observe(got, [{'objective': 0.0}] * 50)                   # observed points stay known
        assert len(dd.history) == 50


def test_primary_algo_with_a_requirement_hands_out_original_points():
    """The design the reference leaves as TODOs (primary_algo.py:13-14,42,55,66):
    the wrapped algorithm works in the transformed space, the caller only ever
    sees points of the original one."""
    from orion_b200.primary_algo import PrimaryAlgo
    from orion_b200.space import Categorical, Integer, Real, Space
    space = Space()
    space.register(Categorical('c', ['a', 'b', 'c', 'd']))
    space.register(Integer('i', 'uniform', -3, 6))
    space.register(Real('x', 'reciprocal', 1e-3, 1.0))
    palgo = PrimaryAlgo(space, {'random': {'seed': 7, 'dedup': True}}, requirement='real')
    inner = palgo.algorithm
    assert inner.space is palgo.transformed_space and palgo.space is space
    assert len(inner.space['c'].shape) == 1 and inner.space['c'].shape[0] == 4      # one-hot inside
    pts = palgo.suggest(6)
    assert len(pts) == 6 and all(p in space for p in pts)
    assert all(p[0] in 'abcd' and int(p[1]) == p[1] for p in pts)
    back = palgo.suggest_columns(500)
    assert not back.transformed and back.n == 500 and bool(back.mask[:500].all())
    assert back.transformed_batch.transformed and back.keep is not None
    palgo.observe(pts, [{'objective': float(j)} for j in range(len(pts))])   # transformed on the way in
    assert palgo.score(pts[0]) == 0 and palgo.judge(pts[0], {}) is None
    before = len(inner.history)
    palgo.observe(pts, [{'objective': 0.0}] * len(pts))                      # already known: nothing new
    assert len(inner.history) == before
    with pytest.raises(AssertionError):
        palgo.observe([('z', 0, 0.5)], [{'objective': 0.0}])
    # without a requirement the wrapper is the snapshot's: same space inside and out
    plain = PrimaryAlgo(space, {'random': {'seed': 7}})
    assert plain.transformed_space is None and plain.algorithm.space is space
    assert all(p in space for p in plain.suggest(3))


def test_suggest_on_a_space_too_wide_for_the_launch_parameters():
    """Up to 192 slots the suggest kernel reads its descriptors from the launch
    parameters (constant bank); wider spaces go through the shared-memory
    staging.  Both must agree with the staged kernels."""
    from orion_b200.space import Categorical, Integer, Real, Space
    from orion_b200.transformer import build_required_space
    for width in (96, 100):                       # 2 * width slots: 192 (parameter path) and 200 (shared memory)
        space = Space()
        space.register(Real('x', 'uniform', -1, 2, shape=width))
        space.register(Integer('k', 'uniform', -5, 10, shape=width // 2))
        space.register(Categorical('c', ['a', 'b', 'c'], shape=width // 2))
        compiled = build_required_space('real', space).compiled()
        assert compiled.n_slots == 2 * width
        n = 1001
        b0, t0, k0 = _staged(compiled, n, 9)
        b1, t1, k1 = compiled.suggest(n, 9)
        _same(b1, b0, n)
        _same(t1, t0, n)
        _same(k1, k0, n)
        assert torch.equal(k1.fp[:n], k0.fp[:n]) and torch.equal(b1.mask[:n], b0.mask[:n])
