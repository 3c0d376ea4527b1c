"""Host logic of the reference-stream mode (orion_b200/mt_stream.py) against the UNMODIFIED reference.

No GPU here: the kernel call (``mt_stream._evaluate``) is replaced by a numpy stand-in that does what
``ob2_sample_from_base`` does per slot, and the compiled space by the few attributes the module reads.
What is under test is everything else -- which RandomState method is asked for how many numbers in
which order, the chunked replay of ``Real.sample``'s four attempts, where the RandomState is left --
by comparing values AND final generator state with the reference's own ``Dimension.sample`` /
``Space.sample`` (src/orion/algo/space.py:110-136, 255-274, 281-294, 394-404, 456-484).
The kernel itself is compared with the reference on the GPU in tests/test_gpu_mt_stream.py.
"""
from collections import OrderedDict

import numpy as np
import pytest

from oracle import ref_harness
from orion_b200 import _lib, mt_stream
from orion_b200 import space as mirror
from orion_b200.descriptor import _draws_by_inverse_cdf, _prior_params

pytestmark = pytest.mark.skipif(not ref_harness.available(), reason='no reference package here')


class HostCompiled(object):
    """The attributes of a CompiledSpace that mt_stream reads, built without a device."""

    def __init__(self, dims):
        self.dims = list(dims)
        self.dim_first, self.dim_widths, self.dim_dists, self.col_meta, self.dim_generic_rvs = [], [], [], [], []
        s = 0
        for d in self.dims:
            k = int(np.prod(d.shape, dtype=np.int64)) if len(d.shape) else 1
            self.dim_first.append(s)
            self.dim_widths.append(k)
            self.dim_dists.append(None if d.type == 'categorical' else _prior_params(d)[0])
            self.dim_generic_rvs.append(d.type != 'categorical' and _draws_by_inverse_cdf(d.prior))
            self.col_meta.extend([(0, {'categorical': 'i4', 'integer': 'i8'}.get(d.type, 'f8'))] * k)
            s += k
        self.n_slots = s
        self.device = 'cpu'


def host_evaluate(compiled, base, n):
    """numpy restatement of one ob2_sample_from_base launch: (columns, row verdicts)."""
    cols, ok = [], np.ones(n, dtype=bool)
    for d, first, k in zip(compiled.dims, compiled.dim_first, compiled.dim_widths):
        for e in range(k):
            b = np.asarray(base[first + e][:n], dtype=np.float64)
            if d.type == 'categorical':
                cdf = np.cumsum(np.asarray(d._probs, dtype=np.float64))
                cdf /= cdf[-1]
                cols.append(np.searchsorted(cdf, b, side='right').astype(np.int32))
                continue
            dist, p0, p1, loc, scale = _prior_params(d)[:5]
            if dist == _lib.DIST_RECIPROCAL:
                b = np.exp(p0 + b * p1)
            elif dist == _lib.DIST_CONST:
                b = np.zeros(n)
            elif dist > _lib.DIST_NONE:             # a further prior: scipy's own inverse CDF stands in for the device's
                args = d.prior._parse_args_rvs(*d._args, size=None, **d._kwargs)[0]
                b = d.prior._ppf(b, *[np.float64(a) for a in args])
            x = b * scale + loc
            low, high = d.interval()
            if d.type == 'integer':
                x = np.floor(x)
            ok &= (x >= low) & (x < high)
            cols.append(x.astype(np.int64) if d.type == 'integer' else x)
    return cols, ok


@pytest.fixture(autouse=True)
def host_kernel(monkeypatch):
    monkeypatch.setattr(mt_stream, '_evaluate', host_evaluate)
    ref_harness.load()
    yield


def ref_classes():
    import orion.algo.space as ref
    return ref


CASES = [
    ('real', ('uniform', -3, 4), {}),
    ('real', ('uniform', -3, 4), {'shape': (2, 3)}),
    ('real', ('norm', 0.9, 0.1), {}),
    ('real', ('norm', 0.9, 0.1), {'shape': 3}),
    ('real', ('reciprocal', 1e-5, 1.0), {}),
    ('real', ('loguniform', 1e-3, 10.0), {'shape': 2}),
    ('real', ('norm', 0.0, 1.0), {'low': -0.5, 'high': 0.7}),                 # about half of the draws rejected
    ('real', ('norm', 0.0, 1.0), {'low': -1.0, 'high': 2.0, 'shape': 2}),     # whole arrays rejected
    ('real', ('norm', 0.0, 1.0), {'low': 0.2, 'high': 2.0, 'shape': 2}),      # ... so often that it gives up
    ('real', ('uniform', 0, 10), {'low': 2.5}),
    ('integer', ('uniform', -3, 6), {}),
    ('integer', ('norm', 5, 3), {'shape': (2, 2)}),
    ('integer', ('uniform', -10, 20), {'low': -2, 'high': 7}),
    ('dimension', ('norm', 0.9, 0.1), {}),
    ('dimension', ('uniform', -3, 4), {'shape': (4, 4)}),
    ('real', ('uniform', 3, 0), {}),                                          # scale 0: a constant, nothing drawn
    ('real', ('fisk', 3.0), {}),                                              # further priors scipy draws as _ppf(uniform())
    ('real', ('genpareto', 0.1, 1.0, 2.0), {'shape': 2}),
    ('integer', ('lomax', 2.0, 0, 10), {'high': 40}),
    ('real', ('invgamma', 3.0), {'low': 0.1, 'high': 2.0}),
    ('real', ('expon', 0.0, 2.0), {}),                                        # numpy generators of their own that are one uniform
    ('real', ('laplace',), {'shape': 2}),                                     # through the inverse CDF
    ('integer', ('logistic', 0, 5), {}),
]


FURTHER = ('fisk', 'genpareto', 'lomax', 'invgamma', 'expon', 'laplace', 'logistic')


def make(module, kind, args, kwargs, name='yolo'):
    klass = {'real': module.Real, 'integer': module.Integer, 'dimension': module.Dimension}[kind]
    return klass(name, *args, **kwargs)


def as_arrays(samples):
    return [np.asarray(s) for s in samples]


def same_state(a, b):
    sa, sb = a.get_state(), b.get_state()
    return sa[0] == sb[0] and np.array_equal(sa[1], sb[1]) and sa[2:] == sb[2:]


@pytest.mark.parametrize('kind,args,kwargs', CASES)
@pytest.mark.parametrize('n', [1, 7, 40])
def test_dimension_sample_matches_the_reference(kind, args, kwargs, n):
    ref_dim, dim = make(ref_classes(), kind, args, kwargs), make(mirror, kind, args, kwargs)
    compiled = HostCompiled([dim])
    for seed in (3, 11):
        ref_rng, rng = np.random.RandomState(seed), np.random.RandomState(seed)
        for _call in range(2):                      # the second call continues the stream
            try:
                expect = as_arrays(ref_dim.sample(n, ref_rng))
            except ValueError as exc:               # improbable bounds: same message, same numbers consumed
                with pytest.raises(ValueError) as mine:
                    mt_stream.sample_dimension(dim, compiled, n, rng)
                assert str(mine.value) == str(exc)
                assert same_state(rng, ref_rng)
                continue
            cols = mt_stream.sample_dimension(dim, compiled, n, rng)
            got = as_arrays([mirror._from_columns(dim, cols, i) for i in range(n)])
            assert len(got) == len(expect) == n
            for g, e in zip(got, expect):
                assert g.shape == e.shape
                if args[0] in FURTHER:              # (numpy's scalar and array pow / expm1 differ in the last digit)
                    assert np.allclose(g, e, rtol=1e-12, atol=1 if kind == 'integer' else 0)
                else:
                    assert np.array_equal(g, e)
            assert same_state(rng, ref_rng)


@pytest.mark.parametrize('kind,args,kwargs', CASES)
def test_integer_seed_reseeds_every_draw(kind, args, kwargs):
    """rvs(random_state=<int>) builds a fresh RandomState per value (space.py:128-131): n copies --
    or the reference's ValueError when that one value misses the bounds."""
    ref_dim, dim = make(ref_classes(), kind, args, kwargs), make(mirror, kind, args, kwargs)
    compiled = HostCompiled([dim])
    for seed in (0, 5, 8):
        try:
            expect = as_arrays(ref_dim.sample(3, seed))
        except ValueError as exc:
            with pytest.raises(ValueError) as mine:
                mt_stream.sample_dimension(dim, compiled, 3, seed)
            assert str(mine.value) == str(exc)
            continue
        cols = mt_stream.sample_dimension(dim, compiled, 3, seed)
        got = as_arrays([mirror._from_columns(dim, cols, i) for i in range(3)])
        for g, e in zip(got, expect):
            assert np.array_equal(g, e) if args[0] not in FURTHER else np.allclose(g, e, rtol=1e-12, atol=0)


def test_global_state_is_the_default_stream():
    ref_dim, dim = make(ref_classes(), 'real', ('norm', 0.0, 1.0), {'low': -1.0}), \
        make(mirror, 'real', ('norm', 0.0, 1.0), {'low': -1.0})
    np.random.seed(5)
    expect = ref_dim.sample(6)
    after = np.random.random_sample()
    np.random.seed(5)
    cols = mt_stream.sample_dimension(dim, HostCompiled([dim]), 6, None)
    assert np.array_equal(cols[0], np.asarray(expect))
    assert np.random.random_sample() == after       # numpy's global generator is left where the reference leaves it


def test_improbable_bounds_raise_after_four_attempts_and_consume_them():
    kwargs = {'low': 8.0, 'high': 9.0}
    ref_dim, dim = make(ref_classes(), 'real', ('norm', 0.0, 1.0), kwargs), make(mirror, 'real', ('norm', 0.0, 1.0), kwargs)
    ref_rng, rng = np.random.RandomState(2), np.random.RandomState(2)
    with pytest.raises(ValueError) as theirs:
        ref_dim.sample(3, ref_rng)
    with pytest.raises(ValueError) as mine:
        mt_stream.sample_dimension(dim, HostCompiled([dim]), 3, rng)
    assert str(mine.value) == str(theirs.value)
    assert same_state(rng, ref_rng)


def test_categorical_matches_the_reference():
    ref = ref_classes()
    for cats, kwargs in ((('asdfa', 2), {}),
                         (OrderedDict(zip(('asdfa', 2, 3, 4), (0.1, 0.2, 0.3, 0.4))), {}),
                         (('a', 'b', 'c'), {'shape': (2, 2)})):
        ref_dim, dim = ref.Categorical('yolo', cats, **kwargs), mirror.Categorical('yolo', cats, **kwargs)
        compiled = HostCompiled([dim])
        for seed in (5, np.random.RandomState(5), None):
            if seed is None:
                np.random.seed(9)
            elif not isinstance(seed, int):
                seed = np.random.RandomState(5)
            expect = ref_dim.sample(9, seed)
            if seed is None:
                np.random.seed(9)
            elif not isinstance(seed, int):
                seed = np.random.RandomState(5)
            cols = mt_stream.sample_dimension(dim, compiled, 9, seed)
            got = [mirror._from_columns(dim, cols, i) for i in range(9)]
            for g, e in zip(got, expect):
                assert np.array_equal(np.asarray(g, dtype=object), np.asarray(e, dtype=object))


def spaces(module):
    space = module.Space()
    space.register(module.Real('lr', 'reciprocal', 1e-5, 1.0))
    space.register(module.Integer('layers', 'uniform', 1, 8))
    space.register(module.Categorical('act', ('relu', 'tanh', 'elu')))
    space.register(module.Real('w', 'norm', 0.0, 1.0, shape=(2,)))
    space.register(module.Real('mom', 'uniform', 0.0, 1.0))
    clipped = module.Space()
    clipped.register(module.Real('a', 'uniform', -1, 2))
    clipped.register(module.Real('b', 'norm', 0.0, 1.0, low=-1.2, high=1.5))        # rejections: the slow path
    clipped.register(module.Categorical('c', OrderedDict((('x', 0.7), ('y', 0.3)))))
    clipped.register(module.Integer('d', 'norm', 3, 2, low=0))
    return [space, clipped]


@pytest.mark.parametrize('which', [0, 1])
@pytest.mark.parametrize('n', [1, 5, 64])
def test_space_sample_matches_the_reference(which, n):
    ref_space, space = spaces(ref_classes())[which], spaces(mirror)[which]
    dims = list(space.values())
    compiled = HostCompiled(dims)
    for seed in (1, 4):
        ref_rng, rng = np.random.RandomState(seed), np.random.RandomState(seed)
        for _call in range(2):
            try:
                expect = ref_space.sample(n, ref_rng)
            except ValueError as exc:
                with pytest.raises(ValueError) as mine:
                    mt_stream.space_columns(dims, compiled, lambda d: HostCompiled([d]), n, rng)
                assert str(mine.value) == str(exc)
                assert same_state(rng, ref_rng)
                continue
            cols = mt_stream.space_columns(dims, compiled, lambda d: HostCompiled([d]), n, rng)
            got = mirror.Space.rows_from_columns(space, cols, n)
            assert len(got) == len(expect) == n
            for grow, erow in zip(got, expect):
                for g, e in zip(grow, erow):
                    assert np.array_equal(np.asarray(g, dtype=object), np.asarray(e, dtype=object))
            assert same_state(rng, ref_rng)


def random_dims(case):
    from tests.test_oracle_vs_reference_random import _random_dims
    return _random_dims(np.random.RandomState(3000 + case))


@pytest.mark.parametrize('case', range(20))
def test_random_spaces_match_the_reference(case):
    """Generated spaces (1-6 dimensions of every kind, shapes, skewed probabilities, user bounds around
    +-2.5 sigma) through the reference and through this module: same points, same final state."""
    from tests import golden_util as gu
    dims_json = random_dims(case)
    ref_space, space = gu.product_space(dims_json, ref_classes()), gu.product_space(dims_json)
    dims = list(space.values())
    compiled = HostCompiled(dims)
    for n in (1, 12):
        ref_rng, rng = np.random.RandomState(case), np.random.RandomState(case)
        try:
            expect = ref_space.sample(n, ref_rng)
        except ValueError as exc:
            with pytest.raises(ValueError) as mine:
                mt_stream.space_columns(dims, compiled, lambda d: HostCompiled([d]), n, rng)
            assert str(mine.value) == str(exc)
            assert same_state(rng, ref_rng)
            continue
        cols = mt_stream.space_columns(dims, compiled, lambda d: HostCompiled([d]), n, rng)
        got = mirror.Space.rows_from_columns(space, cols, n)
        for grow, erow in zip(got, expect):
            for g, e in zip(grow, erow):
                assert np.array_equal(np.asarray(g, dtype=object), np.asarray(e, dtype=object))
        assert same_state(rng, ref_rng)


def test_reference_known_answers_on_this_path():
    """The reference's own golden values (tests/unittests/algo/test_space.py:195-198, 246-250, 266-271)."""
    def draw(dim):
        cols = mt_stream.sample_dimension(dim, HostCompiled([dim]), 1, np.random.RandomState(5))
        return mirror._from_columns(dim, cols, 0)
    assert draw(mirror.Integer('yolo', 'uniform', -3, 6)) == -2
    assert draw(mirror.Categorical('yolo', ('asdfa', 2))) == 'asdfa'
    assert draw(mirror.Categorical('yolo', OrderedDict(zip(('asdfa', 2, 3, 4), (0.1, 0.2, 0.3, 0.4))))) == 2


def test_which_priors_are_reproducible():
    rng = np.random.RandomState(0)
    assert mt_stream.reproducible(HostCompiled([mirror.Real('x', 'uniform', 0, 1)]), rng)
    assert mt_stream.reproducible(HostCompiled([mirror.Real('x', 'expon')]), rng)             # standard_exponential = -log(1 - u)
    assert not mt_stream.reproducible(HostCompiled([mirror.Real('x', 'gamma', 2.0)]), rng)    # standard_gamma
    assert mt_stream.reproducible(HostCompiled([mirror.Real('x', 'fisk', 3.0)]), rng)         # _ppf(uniform())
    assert not mt_stream.reproducible(HostCompiled([mirror.Real('x', 'maxwell')]), rng)       # chi._rvs
    assert not mt_stream.reproducible(HostCompiled([mirror.Real('x', 'uniform', 0, 1)]), np.random.default_rng(0))


def test_numpy_generators_that_are_one_uniform_through_the_inverse_cdf():
    """What mt_stream._UNIFORM_FED relies on, against numpy's legacy RandomState: the same number of uniforms
    consumed, the same values to the last ulp or two (numpy's vectorised log and the C library's differ there)."""
    def same(x, y):
        return np.allclose(x, y, rtol=4.5e-16, atol=0)
    a, b = np.random.RandomState(3), np.random.RandomState(3)
    u = b.random_sample(1000)
    assert same(a.standard_exponential(1000), -np.log(1 - u)) and a.random_sample() == b.random_sample()
    a, b = np.random.RandomState(4), np.random.RandomState(4)
    u = b.random_sample(1000)
    assert same(a.logistic(size=1000), np.log(u / (1 - u))) and a.random_sample() == b.random_sample()
    a, b = np.random.RandomState(5), np.random.RandomState(5)
    u = b.random_sample(1000)
    assert same(a.laplace(0, 1, 1000), np.where(u >= 0.5, -np.log(2 - u - u), np.log(u + u)))
    assert a.random_sample() == b.random_sample()


def test_replay_of_attempts():
    ok = np.array([1, 0, 1, 0, 0, 0, 1, 1], dtype=bool)
    take, used, failed = mt_stream._replay_attempts(ok, 3)
    assert list(take) == [0, 2, 6] and used == 7 and not failed
    take, used, failed = mt_stream._replay_attempts(np.array([1, 0, 0, 0, 0, 1], dtype=bool), 2)
    assert failed and used == 5
    take, used, failed = mt_stream._replay_attempts(np.array([1, 0, 0], dtype=bool), 2)
    assert take is None
