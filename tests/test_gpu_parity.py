"""Parity of the CUDA path (through the C ABI) against the reference's golden
vectors and the numpy oracle.  Bar: bit-exact for Integer / Categorical /
one-hot / reverse / fingerprints / keep masks and for uniform Real; <= 1e-6
relative (fp64) for loguniform and normal."""
import numpy as np
import pytest
import torch

from oracle import space_oracle as so
from tests import golden_util as gu
from tests.spec_util import specs_from_space

pytestmark = pytest.mark.gpu

RTOL = 1e-6


def _pad_U(U):
    S, n = U.shape
    ns = (n + 31) // 32 * 32
    out = np.full((S, ns), 0.5)
    out[:, :n] = U
    return torch.from_numpy(out).cuda(), n


def _compare_cols(specs, got, want):
    table = so.slot_table(specs)
    assert len(got) == len(want) == len(table)
    for s, (d, _e) in enumerate(table):
        spec = specs[d]
        g, w = got[s], want[s]
        assert g.dtype == w.dtype, (s, g.dtype, w.dtype)
        if spec.kind == 'real' and spec.prior in ('reciprocal', 'norm'):
            np.testing.assert_allclose(g, w, rtol=RTOL, atol=0, err_msg='slot %d' % s)
        else:
            assert np.array_equal(g, w), 'slot %d (%s %s)' % (s, spec.kind, spec.prior)


@pytest.mark.parametrize('name', ['mixed_scalar', 'shaped', 'loc_scale_kwargs'])
def test_from_uniform_matches_reference_golden(name):
    g = gu.load(name)
    space = gu.product_space(g['dims'])
    U, n = _pad_U(gu.stream_to_U(g['specs'], g['n'], g['stream']))
    batch = space.compiled().sample_from_uniform(U, n=n)
    _compare_cols(g['specs'], batch.to_numpy(), g['cols'])
    assert bool(batch.mask[:n].all().item())


def test_from_uniform_integer_with_transcendental_prior_is_exact_here():
    """Integer over reciprocal/norm: floor() could flip within a few ulp of an
    integer (SURVEY hard part 2); on the golden streams it must not."""
    g = gu.load('mixed_scalar')
    space = gu.product_space(g['dims'])
    U, n = _pad_U(gu.stream_to_U(g['specs'], g['n'], g['stream']))
    got = space.compiled().sample_from_uniform(U, n=n).to_numpy()
    assert np.array_equal(got[5], g['cols'][5])         # 'li': Integer reciprocal(1, 1000)


@pytest.mark.parametrize('name', gu.CASES)
def test_contains_matches_reference_golden(name):
    from orion_b200.columns import ColumnBatch
    g = gu.load(name)
    space = gu.product_space(g['dims'])
    compiled = space.compiled()
    batch = ColumnBatch.from_numpy(compiled, g['probes'], int_as_f64=True)
    got = compiled.contains(batch, int_as_f64=True).cpu().numpy().astype(bool)
    assert np.array_equal(got, g['probe_verdict'])


@pytest.mark.parametrize('n', [1, 2, 31, 33, 1000, 4097])
def test_philox_matches_oracle_mixed64(n):
    from orion_b200 import spaces
    space = spaces.mixed64()
    specs = specs_from_space(space)
    seed = 0x1234567890ABCDEF
    batch = space.compiled().sample(n, seed, base_index=(1 << 40) + 7, with_fingerprint=True)
    want, failed = so.sample_philox(specs, n, seed, base_index=(1 << 40) + 7)
    assert not failed.any()
    got = batch.to_numpy()
    _compare_cols(specs, got, want)
    assert bool(batch.mask[:n].all().item())
    assert int(batch.n_failed.sum().item()) == 0
    lo, hi = batch.fingerprints_numpy()
    wlo, whi = so.fingerprint(got)
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)


def test_philox_7_rounds_matches_oracle_and_differs_from_10():
    """ob2_space_set_philox_rounds(7): the same plan and arithmetic over Philox4x32-7 (known
    answers: tests/test_oracle_plan.py); transform chain and fingerprints follow."""
    from orion_b200 import spaces
    from orion_b200.transformer import build_required_space
    space = spaces.mixed64()
    specs = specs_from_space(space)
    n, seed = 3001, 424242
    b7 = space.compiled(philox_rounds=7).sample(n, seed, base_index=99, with_fingerprint=True)
    want, failed = so.sample_philox(specs, n, seed, base_index=99, rounds=7)
    assert not failed.any()
    got = b7.to_numpy()
    _compare_cols(specs, got, want)
    b10 = space.compiled().sample(n, seed, base_index=99).to_numpy()
    assert not np.array_equal(got[0], b10[0]) and not np.array_equal(got[40], b10[40])
    lo, hi = b7.fingerprints_numpy()
    wlo, whi = so.fingerprint(got)
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)
    # the one-kernel suggest with 7 rounds hands out the same rows
    sb, st, sk = build_required_space('real', space).compiled(philox_rounds=7).suggest(n, seed, base_index=99)
    for a, b in zip(sk.to_numpy(), got):
        assert np.array_equal(a, b)
    twant = so.transform(specs, 'real', got)
    assert all(np.array_equal(a, b) for a, b in zip(st.to_numpy(), twant))


def test_philox_plan_with_alternating_widths_and_rejections_matches_oracle():
    """Slots of one and two words interleaved (blocks shared across classes, skipped odd words),
    shaped dimensions, a weighted categorical (threshold search), a 100-category one (equal
    probabilities whose cdf is NOT exactly j/100), an Integer that misses the shortcut (user
    bound), a dimension that rejects: device == oracle, value for value."""
    from orion_b200.space import Categorical, Integer, Real, Space
    space = Space()
    space.register(Categorical('c3', {'a': 0.2, 'b': 0.3, 'c': 0.5}))
    space.register(Real('x', 'uniform', 0, 1))
    space.register(Integer('k', 'uniform', 0, 10))
    space.register(Categorical('c100', list(range(100)), shape=(3,)))
    space.register(Real('g', 'norm', 0, 1, low=-1.5, high=1.5, shape=(2,)))      # rejects ~13 % per element
    space.register(Integer('kb', 'uniform', 0, 10, high=7))                       # user bound inside the support
    space.register(Categorical('c2', [True, False]))
    space.register(Real('l', 'reciprocal', 1e-3, 10.0))
    space.register(Integer('kn', 'norm', 0, 5))
    specs = specs_from_space(space)
    plan = so.draw_plan(specs)
    assert [p[2] for p in plan] == [32, 53, 32, 32, 32, 32, 53, 53, 53, 32, 53, 53]
    n, seed = 20000, 5
    batch = space.compiled().sample(n, seed, with_fingerprint=True)
    want, failed = so.sample_philox(specs, n, seed)
    got = batch.to_numpy()
    _compare_cols(specs, got, want)
    mask = batch.mask[:n].cpu().numpy().astype(bool)
    assert np.array_equal(mask, ~failed) and 0 < (~mask).sum() < n // 20        # (four misses in a row: ~1 %)
    first, _ = so.sample_from_uniform(specs, np.stack([so.philox_uniform(seed, np.arange(n), p) for p in plan]))
    assert not np.array_equal(first[6], got[6])                     # some rows of `g` were redrawn
    lo, hi = batch.fingerprints_numpy()
    wlo, whi = so.fingerprint(got)
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)


def test_inverse_cdfs_are_double_precision_grade():
    """The sampler's own ndtri (one logarithm + polynomial) and bounded exp against scipy /
    numpy on a dense sweep of injected uniforms, tails included: far inside the 1e-6 the
    contract allows -- these are not reduced-precision shortcuts."""
    from scipy.special import ndtri
    from orion_b200.space import Real, Space
    space = Space()
    space.register(Real('g', 'norm', 0, 1))
    space.register(Real('l', 'reciprocal', 1e-5, 1.0))
    compiled = space.compiled()
    rng = np.random.RandomState(12)
    u = np.concatenate([rng.random_sample(400000), 10.0 ** rng.uniform(-300, -1, 100000),
                        1 - 10.0 ** rng.uniform(-15.9, -1, 100000), 0.5 + rng.uniform(-1e-7, 1e-7, 20000),
                        np.array([2.0 ** -53, 1 - 2.0 ** -53, 0.5, 5e-324, 1e-310])])
    u = u[(u > 0) & (u < 1)]
    n = len(u)
    U, _ = _pad_U(np.stack([u, u]))
    got = compiled.sample_from_uniform(U, n).to_numpy()
    want_g = ndtri(u)
    nz = want_g != 0
    rel_g = np.abs(got[0][nz] - want_g[nz]) / np.abs(want_g[nz])
    body = u[nz] >= 2.0 ** -53
    assert rel_g[body].max() < 4e-15, rel_g[body].max()
    assert rel_g[~body].max() < 1e-9                                 # below 2^-53: injected uniforms only
    assert got[0][~nz].tolist() == want_g[~nz].tolist()
    want_l = np.exp(np.log(1e-5) + u * (np.log(1.0) - np.log(1e-5)))
    assert (np.abs(got[1] - want_l) / want_l).max() < 1e-15


def test_philox_explicit_counters():
    """index[i] replaces base_index + i: replaying counters reproduces rows."""
    from orion_b200 import spaces
    compiled = spaces.mixed64().compiled()
    seed = 11
    ref = compiled.sample(1000, seed, base_index=500).to_numpy()
    idx = torch.tensor([500 + 7, 500 + 7, 500 + 999, 500, 500 + 123], dtype=torch.int64).cuda()
    got = compiled.sample(5, seed, index=idx, with_fingerprint=True)
    for g, r in zip(got.to_numpy(), ref):
        assert np.array_equal(g, r[[7, 7, 999, 0, 123]])
    lo, hi = got.fingerprints_numpy()
    assert lo[0] == lo[1] and hi[0] == hi[1]


def test_philox_rejection_matches_oracle():
    """Real.sample's <=4 attempts with whole-value redraw, shaped dims included."""
    g = gu.load('bounded_reject')
    space = gu.product_space(g['dims'])
    n, seed = 5000, 77
    batch = space.compiled().sample(n, seed)
    want, failed = so.sample_philox(g['specs'], n, seed)
    got = batch.to_numpy()
    _compare_cols(g['specs'], got, want)
    mask = batch.mask[:n].cpu().numpy().astype(bool)
    assert np.array_equal(mask, ~failed)
    assert int(batch.n_failed.sum().item()) == int(failed.sum())
    # accepted rows are inside the space; rows that used a retry exist
    assert np.array_equal(so.contains(g['specs'], got), ~failed)
    first_try, _ = so.sample_from_uniform(
        g['specs'], np.stack([so.philox_uniform(seed, np.arange(n), s) for s in range(len(got))]))
    assert any(not np.array_equal(a, b) for a, b in zip(first_try, got))


def test_improbable_bounds_raise_like_reference():
    g = gu.load('improbable')
    space = gu.product_space(g['dims'])
    with pytest.raises(ValueError) as exc:
        space.sample(8, seed=3)
    assert str(exc.value) == g['error']
    with pytest.raises(ValueError) as exc:
        space['yolo'].sample(8)
    assert 'Improbable bounds' in str(exc.value)


@pytest.mark.parametrize('requirement', ['real', 'integer', 'numerical', None])
@pytest.mark.parametrize('quantize', ['round', 'floor'])
def test_transform_reverse_match_oracle(requirement, quantize):
    from orion_b200.transformer import build_required_space
    g = gu.load('mixed_scalar')
    dims = g['dims'] + [dict(name='b2', kind='categorical', categories=['x', 'y']),
                        dict(name='b1', kind='categorical', categories=['only']),
                        dict(name='sh', kind='categorical', categories=[1, 2, 3], shape=[2])]
    specs = gu.specs_from_json(dims)
    space = gu.product_space(dims)
    n, seed = 777, 5
    tspace = build_required_space(requirement, space, quantize=quantize)
    compiled = tspace.compiled()
    batch = compiled.sample(n, seed)
    cols = batch.to_numpy()
    tb = compiled.transform(batch)
    tgot = tb.to_numpy()
    twant = so.transform(specs, requirement, cols, quantize=quantize)
    assert len(tgot) == len(twant)
    for a, b in zip(tgot, twant):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    back = compiled.reverse(tb).to_numpy()
    bwant = so.reverse(specs, requirement, twant, quantize=quantize)
    for a, b in zip(back, bwant):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    if requirement in ('real', 'numerical', None):
        for s, (a, b) in enumerate(zip(back, cols)):       # round trip is the identity
            assert np.array_equal(a, b), s


@pytest.mark.parametrize('requirement', ['real', 'integer', 'numerical', None])
def test_fused_sample_transform_equals_two_kernels(requirement):
    """ob2_sample_transform == ob2_sample followed by ob2_transform, bit for
    bit, including rows that went through the rejection path."""
    from orion_b200.transformer import build_required_space
    g = gu.load('bounded_reject')
    extra = [dict(d, name='m_' + d['name']) for d in gu.load('mixed_scalar')['dims'][5:]]
    dims = g['dims'] + extra + [dict(name='b2', kind='categorical', categories=['x', 'y'])]
    space = gu.product_space(dims)
    compiled = build_required_space(requirement, space).compiled()
    n, seed = 3001, 123
    plain = compiled.sample(n, seed, with_fingerprint=True)
    tplain = compiled.transform(plain)
    tfused = compiled.empty_columns(n, transformed=True)
    fused = compiled.sample(n, seed, with_fingerprint=True, transformed_out=tfused)
    for a, b in zip(fused.to_numpy(), plain.to_numpy()):      # (buffers differ only in the stride padding)
        assert a.dtype == b.dtype and np.array_equal(a, b)
    assert torch.equal(fused.fp[:n], plain.fp[:n]) and torch.equal(fused.mask[:n], plain.mask[:n])
    for a, b in zip(tfused.to_numpy(), tplain.to_numpy()):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    assert int(fused.n_failed.sum().item()) == int(plain.n_failed.sum().item())


def test_reverse_argmax_ties_and_soft_values():
    """reverse of arbitrary real vectors: first maximum wins; k == 2 thresholds at 0.5."""
    from orion_b200.columns import ColumnBatch
    from orion_b200.transformer import build_required_space
    dims = [dict(name='c4', kind='categorical', categories=[0, 1, 2, 3]),
            dict(name='c2', kind='categorical', categories=['a', 'b']),
            dict(name='c9', kind='categorical', categories=list(range(9)))]
    specs = gu.specs_from_json(dims)
    space = gu.product_space(dims)
    compiled = build_required_space('real', space).compiled()
    rng = np.random.RandomState(0)
    n = 300
    tcols = [np.round(rng.random_sample(n), 1) for _ in range(4 + 1 + 9)]   # rounding forces ties
    tcols[0][:3], tcols[1][:3], tcols[2][:3], tcols[3][:3] = [.1, .7, .7], [.7, .7, .1], [.7, .1, .7], [0, 0, .7]
    tb = ColumnBatch.from_numpy(compiled, tcols, transformed=True)
    got = compiled.reverse(tb).to_numpy()
    want = so.reverse(specs, 'real', tcols)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)
    assert list(got[0][:3]) == [1, 0, 0]


def test_fingerprint_canonicalises_negative_zero_and_matches_oracle():
    from orion_b200.columns import ColumnBatch
    dims = [dict(name='a', kind='real', prior='uniform', args=[-1, 2]),
            dict(name='i', kind='integer', prior='uniform', args=[-5, 10]),
            dict(name='c', kind='categorical', categories=[0, 1, 2])]
    space = gu.product_space(dims)
    compiled = space.compiled()
    cols = [np.array([0.0, -0.0, 0.5, 0.5]), np.array([1, 1, -2, -2], dtype=np.int64),
            np.array([2, 2, 0, 1], dtype=np.int32)]
    batch = ColumnBatch.from_numpy(compiled, cols)
    compiled.fingerprint(batch)
    lo, hi = batch.fingerprints_numpy()
    wlo, whi = so.fingerprint(cols)
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)
    assert lo[0] == lo[1] and hi[0] == hi[1]
    assert (lo[2], hi[2]) != (lo[3], hi[3])


@pytest.mark.parametrize('n', [3, 1000, 50001])
def test_dedup_matches_exact_oracle(n):
    """Planted duplicates: against history, inside the batch, and both."""
    from orion_b200 import spaces
    from orion_b200.columns import ColumnBatch
    from orion_b200.history import TrialHistory
    space = spaces.mixed64()
    compiled = space.compiled()
    seed = 99
    rng = np.random.RandomState(n)
    hist_cols = compiled.sample(max(n // 2, 1), seed, base_index=10 ** 6).to_numpy()
    fresh = compiled.sample(n, seed).to_numpy()
    # row i: 10 % copy a history row, 5 % copy an earlier batch row (possibly itself a history copy)
    src_hist = rng.randint(0, len(hist_cols[0]), size=n)
    pick_hist = rng.random_sample(n) < 0.10
    cols = [np.where(pick_hist, h[src_hist], f) for h, f in zip(hist_cols, fresh)]
    for i in np.nonzero(rng.random_sample(n) < 0.05)[0]:
        if i > 0:
            j = rng.randint(0, i)
            for c in cols:
                c[i] = c[j]
    oracle = so.ExactHistory()
    oracle.insert(hist_cols)
    want = oracle.dedup(cols)

    hist = TrialHistory(space, capacity=4 * n + 64)
    hist.observe(ColumnBatch.from_numpy(compiled, hist_cols))
    assert len(hist) == len(hist_cols[0])
    batch = ColumnBatch.from_numpy(compiled, cols)
    got = hist.dedup(batch).cpu().numpy().astype(bool)
    assert np.array_equal(got, want)
    assert len(hist) == len(hist_cols[0]) + int(want.sum())
    # idempotence: the same batch again is all duplicates now
    again = hist.dedup(ColumnBatch.from_numpy(compiled, cols)).cpu().numpy()
    assert not again.any()
    # probe agrees with membership
    found = hist.probe_fingerprints(batch.fp, batch.n)[:n].cpu().numpy()
    assert found.all()


def test_dedup_order_keys_and_padding():
    """Explicit order keys (the multi-GPU path): the smallest key wins, and
    entries with fp.lo == 0 are ignored."""
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    space = spaces.rosen2()
    hist = TrialHistory(space, capacity=1024)
    fp = torch.tensor([[5, 1], [5, 1], [7, 2], [0, 0], [5, 1], [9, 3]], dtype=torch.int64).cuda()
    order = torch.tensor([30, 10, 5, 1, 20, 2], dtype=torch.int32).cuda()
    keep = hist.dedup_fingerprints(fp, order=order).cpu().numpy()
    assert list(keep) == [0, 1, 1, 0, 0, 1]
    assert len(hist) == 3


def test_set_full_is_reported():
    from orion_b200 import _lib, spaces
    from orion_b200.history import TrialHistory
    hist = TrialHistory(spaces.rosen2(), n_slots=64)
    fp = torch.arange(1, 2 * 100 + 1, dtype=torch.int64).view(100, 2).cuda()
    with pytest.raises(_lib.Ob2Error) as exc:
        hist.insert_fingerprints(fp)
    assert 'full' in str(exc.value)


def test_empty_and_bad_batches():
    from orion_b200 import _lib, spaces
    compiled = spaces.rosen2().compiled()
    b = compiled.sample(0, 1)
    assert b.n == 0 and b.to_numpy()[0].shape == (0,)
    lib = _lib.load()
    assert lib.ob2_sample(compiled.handle, 5, 7, 0, 0, None, b.buf.data_ptr(), None, None, None, None) == -1
    assert b'n_stride' in lib.ob2_last_error()


@pytest.mark.parametrize('world', [2, 8])
def test_exchange_kernels_emulated_on_one_gpu(world):
    """ob2_bucket_by_owner / ob2_scatter_verdict, and the whole sharded dedup
    with `world` ranks emulated on ONE device: the all_to_all is a transpose of
    the [source][owner] bucket grid, each owner has its own hash set.  Result
    must equal a single set fed the ranks' batches in rank order."""
    from orion_b200 import spaces
    from orion_b200.dist import DeviceOps, bucket_capacity
    from orion_b200.history import TrialHistory
    space = spaces.rosen2()
    ops = DeviceOps('cuda')
    n = 20000
    rng = np.random.RandomState(world)
    pool = rng.randint(1, 2 ** 62, size=(world * n // 2, 2), dtype=np.int64)
    batches = [pool[rng.randint(0, len(pool), size=n)].copy() for _ in range(world)]
    for b in batches:
        b[rng.random_sample(n) < 0.02] = 0                      # padding entries are dropped
    cap = bucket_capacity(n, world)
    grid = []
    for r, b in enumerate(batches):
        out = (torch.full((world, cap, 2), -1, dtype=torch.int64, device='cuda'),
               torch.zeros((world, cap), dtype=torch.int32, device='cuda'),
               torch.zeros((world, cap), dtype=torch.int32, device='cuda'),
               torch.zeros(world, dtype=torch.int32, device='cuda'))
        ops.bucket(torch.from_numpy(b).cuda(), n, r * n, world, cap, out)
        out_fp, out_key, out_src, counts = [t.cpu().numpy() for t in out]
        owner = b[:, 1].view(np.uint64) % np.uint64(world)
        for o in range(world):
            rows = np.nonzero((owner == o) & (b[:, 0] != 0))[0]
            assert counts[o] == len(rows) <= cap
            got = out_src[o, :len(rows)]
            assert sorted(got.tolist()) == rows.tolist()         # every row exactly once, any order
            assert np.array_equal(out_fp[o, :len(rows)], b[got])
            assert np.array_equal(out_key[o, :len(rows)], r * n + got)
            assert not out_fp[o, len(rows):].any()               # zero-padded tail
        grid.append(out)
    # "all_to_all": owner o receives bucket o of every source, in source order
    shards = [TrialHistory(space, capacity=4 * n) for _ in range(world)]
    verdict_back = [[None] * world for _ in range(world)]
    for o in range(world):
        in_fp = torch.stack([grid[r][0][o] for r in range(world)]).view(-1, 2).contiguous()
        in_key = torch.stack([grid[r][1][o] for r in range(world)]).view(-1).contiguous()
        verdict = torch.zeros(world * cap, dtype=torch.uint8, device='cuda')
        ops.dedup(shards[o], in_fp, in_key, verdict)
        for r in range(world):
            verdict_back[r][o] = verdict[r * cap:(r + 1) * cap]
    seen = set()
    for r, b in enumerate(batches):
        keep = torch.zeros(n, dtype=torch.uint8, device='cuda')
        back = torch.cat(verdict_back[r]).contiguous()
        ops.scatter(grid[r][2].view(-1), back, grid[r][0].view(-1, 2), keep)
        want = np.zeros(n, dtype=np.uint8)
        for i, (lo, hi) in enumerate(b):
            if lo != 0 and (int(lo), int(hi)) not in seen:
                seen.add((int(lo), int(hi)))
                want[i] = 1
        assert np.array_equal(keep.cpu().numpy(), want), 'rank %d' % r
    assert sum(len(s) for s in shards) == len(seen)
