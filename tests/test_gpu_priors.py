"""Further scipy priors with a closed-form inverse CDF on the device
(anything `DimensionBuilder._build` accepts,
reference space_builder.py:156-186).  Parity: the device draw from an injected
uniform equals scipy's own ppf within 1e-6 relative; Philox-mode marginals pass
a KS test against scipy's distribution; Integer variants floor exactly."""
import numpy as np
import pytest
import torch
from scipy import stats

from oracle import space_oracle as so
from tests import golden_util as gu

pytestmark = pytest.mark.gpu

PRIORS = [
    ('expon', []), ('expon', [2.0, 3.0]), ('lognorm', [0.5]), ('lognorm', [1.25, 1.0, 2.0]),
    ('logistic', []), ('logistic', [1.0, 0.5]), ('laplace', []), ('gumbel_r', [0.5, 2.0]),
    ('weibull_min', [1.5]), ('weibull_min', [0.7, 1.0, 3.0]), ('pareto', [2.5]), ('rayleigh', []),
    ('rayleigh', [0.0, 4.0]), ('halfnorm', []), ('powerlaw', [3.0]), ('powerlaw', [0.5, -1.0, 2.0]),
    ('gamma', [0.3]), ('gamma', [2.5, 1.0, 3.0]), ('gamma', [400.0]), ('erlang', [3]),
    ('beta', [2.0, 5.0]), ('beta', [0.5, 0.5]), ('beta', [30.0, 2.0, -1.0, 4.0]), ('beta', [1.0, 3.0]),
    ('cauchy', []), ('cauchy', [1.0, 0.1]), ('chi2', [3.0]), ('chi2', [0.7, 0.0, 2.0]),
    ('t', [1.0]), ('t', [2.5]), ('t', [30.0, 1.0, 0.5]),
    ('truncnorm', [-1.0, 2.0]), ('truncnorm', [0.5, 3.0, 10.0, 2.0]), ('truncnorm', [-3.0, -1.0]), ('truncnorm', [2.0, 6.0]),
]


def _space(kind='real'):
    dims = [dict(name='d%d' % j, kind=kind, prior=p, args=a) for j, (p, a) in enumerate(PRIORS)]
    return dims, gu.specs_from_json(dims), gu.product_space(dims)


def test_from_uniform_matches_scipy_ppf():
    dims, specs, space = _space()
    rng = np.random.RandomState(0)
    n = 4096
    U = rng.random_sample((len(dims), n))
    U[:, :4] = [1e-300, 2.0 ** -53, 0.5, 1 - 2.0 ** -53]
    for j, (name, args) in enumerate(PRIORS):
        if (name == 'gamma' and args[0] > 100) or name == 't':
            U[j, 0] = 1e-100       # P(400, x) / the beta variable of t underflow below ~1e-250 (a Philox u is 0 or >= 2^-53)
        if name == 'truncnorm':
            U[j, 0] = 1e-10        # at u ~ 0 the draw IS the bound a: in or out of [a, b) by one ulp of ndtri
    batch = space.compiled().sample_from_uniform(torch.from_numpy(U).cuda())
    got = batch.to_numpy()
    want, ok = so.sample_from_uniform(specs, U)
    for j, (g, w) in enumerate(zip(got, want)):
        np.testing.assert_allclose(g, w, rtol=1e-6, atol=1e-13, err_msg='%s%r' % PRIORS[j])
        direct = getattr(stats, PRIORS[j][0]).ppf(U[j], *PRIORS[j][1])
        np.testing.assert_allclose(g, direct, rtol=1e-6, atol=1e-13, err_msg='%s%r' % PRIORS[j])
    assert np.array_equal(batch.mask.cpu().numpy().astype(bool), ok)


def test_philox_marginals_pass_ks():
    dims, specs, space = _space()
    n = 400_000
    cols = space.compiled().sample(n, 31337).to_numpy()
    for j, x in enumerate(cols):
        dist = getattr(stats, PRIORS[j][0])(*PRIORS[j][1])
        p = stats.kstest(x, dist.cdf).pvalue
        assert p > 1e-3 / len(cols), '%s%r: p = %g' % (PRIORS[j] + (p,))


def test_integer_variants_and_bounds():
    """discrete=True over these priors: floor of the continuous draw, checked
    against the floored interval (reference space.py:281-319)."""
    dims, specs, space = _space('integer')
    n = 50_000
    batch = space.compiled().sample(n, 5)
    got = batch.to_numpy()
    want, failed = so.sample_philox(specs, n, 5)
    flips = 0
    for j, (g, w) in enumerate(zip(got, want)):
        assert g.dtype == np.int64
        flips += int((g != w).sum())        # floor can flip only when the draw is within ulps of an integer
    assert flips <= 2
    assert bool(batch.mask[:n].all().item()) and not failed.any()
    assert np.array_equal(so.contains(specs, got), np.ones(n, dtype=bool))


DSL_EXAMPLES = [      # the priors of the reference's own DSL examples (space_builder.py:84,98-99)
    dict(name='a', kind='real', prior='alpha', args=[0.9]),
    dict(name='a2', kind='real', prior='alpha', args=[2.5, 1.0, 3.0]),
    dict(name='r', kind='integer', prior='randint', args=[-3, 5]),
    dict(name='r2', kind='integer', prior='randint', args=[0, 1000], kwargs=dict(loc=7)),
    dict(name='p', kind='integer', prior='poisson', args=[], kwargs=dict(mu=3)),
    dict(name='p2', kind='integer', prior='poisson', args=[0.5], kwargs=dict(loc=2)),
    dict(name='p3', kind='integer', prior='poisson', args=[120.0]),
    dict(name='gm', kind='integer', prior='geom', args=[0.3]),
    dict(name='gm2', kind='integer', prior='geom', args=[0.02], kwargs=dict(loc=-1)),
    dict(name='bn', kind='integer', prior='binom', args=[20, 0.3]),
    dict(name='bn2', kind='integer', prior='binom', args=[100, 0.9]),
]


def test_dsl_example_priors_from_uniform():
    """alpha within 1e-6 of scipy's ppf; randint (scipy's _ppf operation for
    operation) exact; poisson exact up to pmf-summation ulps at a cdf step."""
    specs = gu.specs_from_json(DSL_EXAMPLES)
    space = gu.product_space(DSL_EXAMPLES)
    rng = np.random.RandomState(1)
    n = 8192
    U = rng.random_sample((len(specs), n))
    U[:, :4] = [0.0, 2.0 ** -53, 0.5, 1 - 2.0 ** -53]
    U[0:2, 0] = 1e-300                                        # alpha.ppf(0) = 1/inf
    U[0:2, 3] = 1 - 1e-9      # (beyond that scipy's own formula 1/(a - ndtri(q ndtr(a))) cancels catastrophically)
    batch = space.compiled().sample_from_uniform(torch.from_numpy(U).cuda())
    got = batch.to_numpy()
    want, ok = so.sample_from_uniform(specs, U)
    for j in (0, 1):
        np.testing.assert_allclose(got[j], want[j], rtol=1e-6, atol=1e-300)
        np.testing.assert_allclose(got[j], stats.alpha.ppf(U[j], *DSL_EXAMPLES[j]['args']), rtol=1e-6, atol=1e-300)
    for j in (2, 3):
        assert got[j].dtype == np.int64 and np.array_equal(got[j], want[j])
    for j in (4, 5, 6, 7, 8, 9, 10):
        assert int((got[j] != want[j]).sum()) <= 1
    assert int((batch.mask.cpu().numpy().astype(bool) != ok).sum()) <= 1


def test_dsl_example_priors_philox_marginals():
    """Philox mode: the accepted draws follow scipy's distribution conditioned
    on the dimension's interval (randint(-3, 5) -> [-4, 4): the reference's own
    interval excludes the top value, space.py:296-319)."""
    specs = gu.specs_from_json(DSL_EXAMPLES)
    space = gu.product_space(DSL_EXAMPLES)
    n = 400_000
    batch = space.compiled().sample(n, 2024)
    ok = batch.mask[:n].cpu().numpy().astype(bool)
    cols = batch.to_numpy()
    want, failed = so.sample_philox(specs, n, 2024)
    assert np.array_equal(ok, ~failed)
    for j, spec in enumerate(DSL_EXAMPLES):
        args, kw = spec['args'], spec.get('kwargs', {})
        dist = getattr(stats, spec['prior'])(*args, **kw)
        x = cols[j][ok]
        if spec['kind'] == 'real':
            p = stats.kstest(x, dist.cdf).pvalue
        else:
            lo, hi = so.dim_interval(specs[j])
            top = min(hi, dist.ppf(1 - 1e-12) + 1)
            ks = np.arange(max(lo, dist.ppf(0) + 1 if np.isinf(lo) else lo), top)
            pmf = dist.pmf(ks)
            ks, pmf = ks[pmf > 0], pmf[pmf > 0]
            exp = pmf / pmf.sum() * len(x)                              # conditioned on the interval
            obs = np.array([(x == k).sum() for k in ks], dtype=np.float64)
            assert obs.sum() == len(x)                                   # nothing outside the interval
            big = exp >= 5                                               # pool the sparse tail
            if (~big).any():
                obs = np.append(obs[big], obs[~big].sum())
                exp = np.append(exp[big], exp[~big].sum())
            p = stats.chisquare(obs, exp).pvalue
            assert int((cols[j] != want[j]).sum()) <= 1
        assert p > 1e-3 / len(DSL_EXAMPLES), '%s: p = %g' % (spec['name'], p)


def test_unsupported_prior_still_refuses():
    from orion_b200.space import Real, Space
    space = Space()
    space.register(Real('a', 'ncx2', 3.0, 1.0))
    assert (0.5,) in space
    with pytest.raises(NotImplementedError):
        space.sample(1)


# ---- the second batch of closed forms (ob2.h OB2_DIST_GUMBEL_L .. OB2_DIST_WRAPCAUCHY)
PRIORS_2 = [
    ('gumbel_l', []), ('gumbel_l', [1.0, 2.0]), ('genpareto', [0.1]), ('genpareto', [-0.1]), ('genpareto', [0.0]),
    ('genpareto', [0.5, 1.0, 2.0]), ('lomax', [2.0]), ('burr', [2.0, 1.5]), ('fisk', [3.0]), ('burr12', [2.0, 1.5]),
    ('invweibull', [2.5]), ('truncexpon', [4.0]), ('bradford', [2.0]), ('arcsine', []), ('arcsine', [-1.0, 2.0]),
    ('triang', [0.3]), ('triang', [0.5, -1.0, 2.0]), ('loglaplace', [2.0]), ('exponpow', [1.5]), ('gompertz', [0.7]),
    ('genlogistic', [0.8]), ('hypsecant', []), ('halfcauchy', []), ('halflogistic', []), ('chi', [3.5]), ('maxwell', []),
    ('maxwell', [0.0, 2.0]), ('invgamma', [3.0]), ('nakagami', [1.7]), ('gibrat', []), ('genextreme', [0.2]),
    ('genextreme', [-0.3]), ('genextreme', [0.0]), ('kappa3', [1.5]), ('dweibull', [1.3]), ('anglit', []),
    ('exponweib', [2.0, 1.5]), ('mielke', [2.0, 3.0]), ('genhalflogistic', [0.7]), ('tukeylambda', [0.3]),
    ('tukeylambda', [0.0]), ('levy', []), ('levy_l', []), ('weibull_max', [1.8]), ('rdist', [4.0]), ('semicircular', []),
    ('dgamma', [1.5]), ('gennorm', [1.5]), ('gennorm', [0.7, 1.0, 2.0]), ('moyal', []), ('wrapcauchy', [0.4]),
]


def _space2(kind='real'):
    dims = [dict(name='e%d' % j, kind=kind, prior=p, args=a) for j, (p, a) in enumerate(PRIORS_2)]
    return dims, gu.specs_from_json(dims), gu.product_space(dims)


def test_second_batch_from_uniform_matches_scipy_ppf():
    """Random uniforms plus 1e-12, 1e-6, 1/2 and 1 - 1e-6 (closer to 1 several of scipy's own formulas cancel:
    burr's q^(-1/d) - 1, kappa3, mielke ...)."""
    dims, specs, space = _space2()
    rng = np.random.RandomState(2)
    n = 4096
    U = rng.random_sample((len(dims), n))
    U[:, :4] = [1e-12, 1e-6, 0.5, 1 - 1e-6]
    batch = space.compiled().sample_from_uniform(torch.from_numpy(U).cuda())
    got = batch.to_numpy()
    want, ok = so.sample_from_uniform(specs, U)
    for j, (g, w) in enumerate(zip(got, want)):
        np.testing.assert_allclose(g, w, rtol=1e-6, atol=1e-12, err_msg='%s%r' % PRIORS_2[j])
        direct = getattr(stats, PRIORS_2[j][0]).ppf(U[j], *PRIORS_2[j][1])
        np.testing.assert_allclose(g, direct, rtol=1e-6, atol=1e-12, err_msg='%s%r' % PRIORS_2[j])
    assert int((batch.mask.cpu().numpy().astype(bool) != ok).sum()) <= 2       # a draw within an ulp of a bound


def test_second_batch_philox_marginals_pass_ks():
    dims, specs, space = _space2()
    n = 200_000
    batch = space.compiled().sample(n, 4242)
    cols = batch.to_numpy()
    for j, x in enumerate(cols):
        dist = getattr(stats, PRIORS_2[j][0])(*PRIORS_2[j][1])
        p = stats.kstest(x, dist.cdf).pvalue
        assert p > 1e-3 / len(cols), '%s%r: p = %g' % (PRIORS_2[j] + (p,))
    assert np.array_equal(so.contains(specs, cols), np.ones(n, dtype=bool))
    assert int(batch.n_failed.sum().item()) == 0


def test_second_batch_through_the_reference_dsl_names():
    """Any scipy name DimensionBuilder._build accepts (space_builder.py:173-186) compiles; the ones without a
    device inverse CDF still refuse to sample."""
    from orion_b200.space import Real, Space
    space = Space()
    space.register(Real('a', 'fisk', 3.0))
    space.register(Real('b', 'ncx2', 3.0, 1.0))
    with pytest.raises(NotImplementedError):
        space.sample(1)
    ok = Space()
    ok.register(Real('a', 'fisk', 3.0))
    ok.register(Real('m', 'maxwell', low=0.5, high=3.0))
    pts = ok.sample(50, seed=3)
    assert len(pts) == 50 and all(p in ok for p in pts) and all(0.5 <= p[1] < 3.0 for p in pts)


# ---- the priors VERDICT r1 listed by name: f, loggamma (closed over the beta / gamma inverses), nbinom, hypergeom
# (pmf summation), vonmises (Newton on the cdf series scipy sums; scipy itself inverts numerically)
PRIORS_3 = [
    dict(name='f', kind='real', prior='f', args=[5.0, 7.0]),
    dict(name='f2', kind='real', prior='f', args=[100.0, 1.5, 1.0, 2.0]),
    dict(name='f3', kind='real', prior='f', args=[1.0, 30.0]),
    dict(name='lg', kind='real', prior='loggamma', args=[0.5]),
    dict(name='lg2', kind='real', prior='loggamma', args=[3.0, 1.0, 2.0]),
    dict(name='vm', kind='real', prior='vonmises', args=[2.0]),
    dict(name='vm2', kind='real', prior='vonmises', args=[0.3, 1.0, 0.5]),
    dict(name='vm3', kind='real', prior='vonmises', args=[30.0]),
    dict(name='nb', kind='integer', prior='nbinom', args=[5, 0.3]),
    dict(name='nb2', kind='integer', prior='nbinom', args=[2.5, 0.9], kwargs=dict(loc=3)),
    dict(name='nb3', kind='integer', prior='nbinom', args=[40, 0.05]),
    dict(name='hg', kind='integer', prior='hypergeom', args=[50, 10, 20]),
    dict(name='hg2', kind='integer', prior='hypergeom', args=[30, 25, 12]),
    dict(name='hg3', kind='integer', prior='hypergeom', args=[5000, 300, 1000]),
]


def test_named_priors_from_uniform_match_scipy():
    specs = gu.specs_from_json(PRIORS_3)
    space = gu.product_space(PRIORS_3)
    rng = np.random.RandomState(4)
    n = 512                                                   # (scipy's numerical vonmises.ppf takes ~2 ms a value)
    U = rng.random_sample((len(specs), n))
    U[:, :4] = [1e-9, 1e-4, 0.5, 1 - 1e-6]
    batch = space.compiled().sample_from_uniform(torch.from_numpy(U).cuda())
    got = batch.to_numpy()
    want, ok = so.sample_from_uniform(specs, U)
    for j, d in enumerate(PRIORS_3):
        if d['kind'] == 'real':
            np.testing.assert_allclose(got[j], want[j], rtol=1e-6, atol=1e-9, err_msg=d['name'])
        else:
            assert got[j].dtype == np.int64 and int((got[j] != want[j]).sum()) <= 1, d['name']
    assert int((batch.mask.cpu().numpy().astype(bool) != ok).sum()) <= 2


def test_named_priors_philox_marginals():
    space = gu.product_space(PRIORS_3)
    n = 200_000
    cols = space.compiled().sample(n, 777).to_numpy()
    for d, x in zip(PRIORS_3, cols):
        dist = getattr(stats, d['prior'])(*d['args'], **d.get('kwargs', {}))
        if d['kind'] == 'real':
            p = stats.kstest(x, dist.cdf).pvalue
            assert p > 1e-4, '%s: p = %g' % (d['name'], p)
        else:                                                  # frequencies against the pmf (only in-bounds values survive:
            lo, hi = space[d['name']].interval()               # the reference's half-open interval drops the top value;
            x = x[(x >= lo) & (x < hi)]                        # a row that missed four times keeps its last, flagged, value)
            assert len(x) > n - 50
            ks, counts = np.unique(x, return_counts=True)
            inside = np.arange(max(lo, dist.support()[0]), min(hi, ks.max() + 1))
            mass = dist.pmf(inside).sum()
            for k, c in zip(ks, counts):
                expect = dist.pmf(k) / mass
                assert abs(c / len(x) - expect) < 5 * np.sqrt(expect / n) + 1e-4, (d['name'], k, c / len(x), expect)


# ---- a third batch of closed forms (ob2.h OB2_DIST_JOHNSONSB .. OB2_DIST_BOLTZMANN)
PRIORS_4 = [
    dict(name='jb', kind='real', prior='johnsonsb', args=[0.5, 1.5]),
    dict(name='ju', kind='real', prior='johnsonsu', args=[-0.5, 2.0, 1.0, 3.0]),
    dict(name='fl', kind='real', prior='fatiguelife', args=[0.7]),
    dict(name='pn', kind='real', prior='powernorm', args=[2.5]),
    dict(name='pl', kind='real', prior='powerlognorm', args=[2.0, 0.5]),
    dict(name='gg', kind='real', prior='gengamma', args=[2.0, 1.5]),
    dict(name='gg2', kind='real', prior='gengamma', args=[2.0, -1.5]),
    dict(name='k4', kind='real', prior='kappa4', args=[0.5, 0.3]),
    dict(name='k4b', kind='real', prior='kappa4', args=[0.0, 0.3]),
    dict(name='k4c', kind='real', prior='kappa4', args=[0.5, 0.0]),
    dict(name='k4d', kind='real', prior='kappa4', args=[0.0, 0.0]),
    dict(name='la', kind='real', prior='laplace_asymmetric', args=[1.7]),
    dict(name='tz', kind='real', prior='trapezoid', args=[0.2, 0.7]),
    dict(name='tp', kind='real', prior='truncpareto', args=[2.0, 5.0]),
    dict(name='tw', kind='real', prior='truncweibull_min', args=[1.5, 0.5, 3.0]),
    dict(name='sc', kind='real', prior='skewcauchy', args=[0.3]),
    dict(name='pk', kind='integer', prior='planck', args=[0.5]),
    dict(name='dl', kind='integer', prior='dlaplace', args=[0.8]),
    dict(name='bz', kind='integer', prior='boltzmann', args=[1.4, 19]),
    dict(name='be', kind='integer', prior='bernoulli', args=[0.3]),       # (the half-open interval leaves only 0)
]


def test_third_batch_from_uniform_matches_scipy():
    specs = gu.specs_from_json(PRIORS_4)
    space = gu.product_space(PRIORS_4)
    rng = np.random.RandomState(6)
    n = 4096
    U = rng.random_sample((len(specs), n))
    U[:, :4] = [1e-9, 1e-4, 0.5, 1 - 1e-6]
    batch = space.compiled().sample_from_uniform(torch.from_numpy(U).cuda())
    got = batch.to_numpy()
    want, ok = so.sample_from_uniform(specs, U)
    for j, d in enumerate(PRIORS_4):
        if d['kind'] == 'real':
            np.testing.assert_allclose(got[j], want[j], rtol=1e-6, atol=1e-12, err_msg=d['name'])
        else:
            assert got[j].dtype == np.int64 and int((got[j] != want[j]).sum()) <= 1, d['name']
    assert int((batch.mask.cpu().numpy().astype(bool) != ok).sum()) <= 2


def test_third_batch_philox_marginals():
    dims = PRIORS_4[:-1]
    space = gu.product_space(dims)
    n = 200_000
    cols = space.compiled().sample(n, 999).to_numpy()
    for d, x in zip(dims, cols):
        dist = getattr(stats, d['prior'])(*d['args'])
        if d['kind'] == 'real':
            p = stats.kstest(x, dist.cdf).pvalue
            assert p > 1e-4, '%s: p = %g' % (d['name'], p)
        else:
            lo, hi = space[d['name']].interval()
            x = x[(x >= lo) & (x < hi)]
            assert len(x) > n - 50
            ks, counts = np.unique(x, return_counts=True)
            inside = np.arange(max(lo, max(dist.support()[0], ks.min())), min(hi, ks.max() + 1))
            mass = dist.pmf(inside).sum()
            for k, c in zip(ks, counts):
                expect = dist.pmf(k) / mass
                assert abs(c / len(x) - expect) < 5 * np.sqrt(expect / n) + 2e-4, (d['name'], k, c / len(x), expect)
