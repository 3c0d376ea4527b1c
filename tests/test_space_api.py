"""The mirror of ``orion.algo.space`` behaves like the reference's classes.

These follow the reference's own ``tests/unittests/algo/test_space.py`` case
by case (same constructions, same expected intervals / membership / errors /
reprs), INCLUDING its equalities with MT19937-seeded scipy draws: the
list-returning calls feed the kernel the reference's own random stream
(orion_b200/mt_stream.py), so ``dists.norm.rvs(0.9, 0.1) == samples[0]`` and the
known answers -2 / 'asdfa' / 2 hold as written there.  Construction-only checks
run on CPU; anything that samples or tests membership needs the GPU.
"""
from collections import OrderedDict

import numpy as np
import pytest
from scipy.stats import distributions as dists

from orion_b200.space import Categorical, Dimension, Integer, Real, Space

gpu = pytest.mark.gpu


@pytest.fixture(scope='function')
def seed():
    """Fixed RandomState and global seed (reference test_space.py:15-21)."""
    rng = np.random.RandomState(5)
    np.random.seed(5)
    return rng


class TestDimension(object):

    @gpu
    def test_simple_instance(self, seed):
        dim = Dimension('yolo', 'norm', 0.9, 0.1)
        samples = dim.sample(seed=seed)
        assert len(samples) == 1
        assert isinstance(samples[0], np.float64)
        assert dists.norm.rvs(0.9, 0.1) == samples[0]           # test_space.py:30
        assert dists.norm.interval(1.0, 0.9, 0.1) == dim.interval()
        assert dists.norm.interval(0.5, 0.9, 0.1) == dim.interval(0.5)
        assert 1.0 in dim
        assert str(dim) == "Dimension(name=yolo, prior={norm: (0.9, 0.1), {}}, shape=())"
        assert dim.name == 'yolo'
        assert dim.type == 'dimension'
        assert dim.shape == ()

    @gpu
    def test_shaped_instance(self, seed):
        dim = Dimension('yolo', 'norm', 0.9, shape=(3, 2))
        samples = dim.sample(seed=seed)
        assert len(samples) == 1
        assert samples[0].shape == (3, 2) and samples[0].dtype == np.float64
        assert np.array_equal(dists.norm.rvs(0.9, size=(3, 2)), samples[0])     # test_space.py:48
        assert dim.shape == (3, 2)
        dim = Dimension('yolo', 'norm', 0.9, shape=4)
        samples = dim.sample(seed=seed)
        assert samples[0].shape == (4,)
        assert np.array_equal(dists.norm.rvs(0.9, size=4), samples[0])          # test_space.py:55
        assert dim.shape == (4,)

    def test_ban_size_kwarg(self):
        with pytest.raises(ValueError):
            Dimension('yolo', 'norm', 0.9, size=(3, 2))

    def test_ban_seed_kwarg(self):
        with pytest.raises(ValueError):
            Dimension('yolo', 'norm', 0.9, seed=8)

    def test_ban_rng_kwarg(self):
        with pytest.raises(ValueError):
            Dimension('yolo', 'norm', 0.9, random_state=8)

    @gpu
    def test_with_predefined_dist(self, seed):
        dim = Dimension('yolo', dists.norm, 0.9)
        samples = dim.sample(seed=seed)
        assert len(samples) == 1
        assert dists.norm.rvs(0.9) == samples[0]                # test_space.py:79

    def test_ban_discrete_kwarg(self):
        with pytest.raises(ValueError) as exc:
            Dimension('yolo', 'uniform', -3, 4, shape=(4, 4), discrete=True)
        assert "pure `_Discrete`" in str(exc.value)

    @gpu
    def test_many_samples(self, seed):
        dim = Dimension('yolo', 'uniform', -3, 4, shape=(4, 4))
        samples = dim.sample(n_samples=4, seed=seed)
        assert len(samples) == 4
        assert np.array_equal(dists.uniform.rvs(-3, 4, size=(4, 4)), samples[0])    # test_space.py:93
        assert all(s.shape == (4, 4) for s in samples)
        assert not np.array_equal(samples[0], samples[1])
        assert all(s in dim for s in samples)

    @gpu
    def test_int_seed_repeats_and_random_state_advances(self):
        """space.py:118-131: an int seed gives the same draw every call -- every VALUE, since
        scipy seeds afresh per rvs; a RandomState carries on."""
        dim = Real('yolo', 'uniform', -3, 4)
        assert dim.sample(3, seed=7) == dim.sample(3, seed=7)
        assert dim.sample(3, seed=7) == [dists.uniform.rvs(-3, 4, random_state=7)] * 3
        rng = np.random.RandomState(7)
        assert dim.sample(3, seed=rng) != dim.sample(3, seed=rng)

    def test_interval(self):
        dim = Dimension('yolo', 'uniform', -3, 4)
        assert dim.interval(1.0) == (-3.0, 1.0)  # `scale` is not the upper bound

    @gpu
    def test_contains_bounds(self):
        dim = Dimension('yolo', 'uniform', -3, 4)
        assert -3 in dim
        assert -2 in dim
        assert 0 in dim
        assert 0.9999 in dim
        assert 1 not in dim

    @gpu
    def test_contains_shape(self):
        dim = Dimension(None, 'uniform', -3, 4, shape=(4, 4))
        assert -3 not in dim
        assert -2 not in dim
        assert 0 not in dim
        assert 0.9999 not in dim
        assert 1 not in dim
        assert dists.uniform.rvs(-3, 4, size=(4, 4)) in dim

    def test_set_bad_name(self):
        dim = Dimension('yolo', 'uniform', -3, 4, shape=(4, 4))
        with pytest.raises(TypeError):
            dim.name = 4


class TestReal(object):

    @gpu
    def test_simple_instance(self, seed):
        dim = Real('yolo', 'norm', 0.9)
        samples = dim.sample(seed=seed)
        assert len(samples) == 1
        assert dists.norm.rvs(0.9) == samples[0]                # test_space.py:134
        assert dists.norm.interval(1.0, 0.9) == dim.interval()
        assert dists.norm.interval(0.5, 0.9) == dim.interval(0.5)
        assert 1.0 in dim
        assert str(dim) == "Real(name=yolo, prior={norm: (0.9,), {}}, shape=())"
        assert dim.name == 'yolo'
        assert dim.type == 'real'
        assert dim.shape == ()

    @gpu
    def test_contains_extra_bounds(self):
        dim = Real('yolo', 'norm', 0, 3, low=-3, high=+3)
        assert dists.uniform.rvs(-3, 3) in dim
        assert -4 not in dim
        assert +4 not in dim
        assert (1, 2) not in dim

    @gpu
    def test_sample_from_extra_bounds_good(self):
        dim = Real('yolo', 'norm', 0, 2, low=-5, high=+5, shape=(4, 4))
        for _ in range(8):
            samples = dim.sample(8)
            for sample in samples:
                assert sample in dim

    @gpu
    def test_sample_from_extra_bounds_bad(self):
        dim = Real('yolo', 'norm', 0, 2, low=-2, high=+2, shape=(4, 4))
        with pytest.raises(ValueError) as exc:
            dim.sample(8)
        assert "Improbable bounds" in str(exc.value)

    def test_bad_bounds(self):
        with pytest.raises(ValueError):
            Real('yolo', 'norm', 0, 2, low=+2, high=-2, shape=(4, 4))
        with pytest.raises(ValueError):
            Real('yolo', 'norm', 0, 2, low=+2, high=+2, shape=(4, 4))

    def test_interval(self):
        dim = Real('yolo', 'norm', 0, 3, low=-3, high=+3)
        assert dim.interval() == (-3, 3)
        dim = Real('yolo', 'alpha', 0.9, low=-3, high=+3)
        assert dim.interval() == (0, 3)
        dim = Real('yolo', 'uniform', -2, 4, low=-3, high=+3)
        assert dim.interval() == (-2.0, 2.0)


class TestInteger(object):

    @gpu
    def test_simple_instance(self, seed):
        dim = Integer('yolo', 'uniform', -3, 6)
        samples = dim.sample(seed=seed)
        assert len(samples) == 1
        assert isinstance(samples[0], np.int64)
        assert samples[0] == -2                                 # test_space.py:198
        assert dim.interval() == (-3, 3)
        assert dim.interval(0.5) == (-2, 2)
        assert 1.0 in dim
        assert str(dim) == "Integer(name=yolo, prior={uniform: (-3, 6), {}}, shape=())"
        assert dim.name == 'yolo'
        assert dim.type == 'integer'
        assert dim.shape == ()

    @gpu
    def test_contains(self):
        dim = Integer('yolo', 'uniform', -3, 6)
        assert 0.1 not in dim
        assert (0.1, -0.2) not in dim
        assert 0 in dim
        assert (1, 2) not in dim
        assert 6 not in dim
        assert -3 in dim
        assert -4 not in dim

    def test_interval_with_infs(self):
        dim = Integer('yolo', 'poisson', 5)
        assert dim.interval()[1] == np.inf

    @gpu
    def test_unsupported_prior_has_no_cpu_fallback(self):
        dim = Integer('yolo', 'skellam', 3.0, 2.0)                       # no device inverse CDF
        assert 3 in dim
        with pytest.raises(NotImplementedError):
            dim.sample()
        draws = Integer('yolo', 'hypergeom', 20, 7, 12).sample(50, seed=1)
        assert all(int(x) == x and 0 <= x <= 7 for x in draws) and len(set(draws)) > 2
        draws = Integer('yolo', 'poisson', 5).sample(50, seed=1)         # this one has a device inverse CDF
        assert all(int(x) == x and x >= 0 for x in draws) and len(set(draws)) > 3


class TestCategorical(object):

    @gpu
    def test_with_tuple(self, seed):
        categories = ('asdfa', 2)
        dim = Categorical('yolo', categories)
        samples = dim.sample(seed=seed)
        assert len(samples) == 1
        assert samples[0] == 'asdfa'                            # test_space.py:250
        assert dim._probs == (0.5, 0.5)
        assert categories == dim.categories
        assert 2 in dim
        assert 3 not in dim
        assert str(dim) == "Categorical(name=yolo, prior={asdfa: 0.50, 2: 0.50}, shape=())"
        assert dim.name == 'yolo'
        assert dim.type == 'categorical'
        assert dim.shape == ()

    @gpu
    def test_with_dict(self, seed):
        probs = (0.1, 0.2, 0.3, 0.4)
        categories = ('asdfa', 2, 3, 4)
        dim = Categorical('yolo', OrderedDict(zip(categories, probs)))
        samples = dim.sample(seed=seed)
        assert len(samples) == 1
        assert samples[0] == 2                                  # test_space.py:271
        assert dim._probs == probs
        assert categories == dim.categories
        assert 2 in dim
        assert 0 not in dim

    @gpu
    def test_probabilities_are_ok(self, seed):
        """Law of large numbers on the category frequencies (test_space.py:283-296)."""
        probs = (0.1, 0.2, 0.3, 0.4)
        categories = OrderedDict(zip(('asdfa', '2', '3', '4'), probs))
        dim = Categorical('yolo', categories)
        samples = dim.sample(50000, seed=seed)
        for key, value in categories.items():
            assert abs(samples.count(key) / 50000. - value) < 0.01

    @gpu
    def test_contains_wrong_shape(self):
        categories = {'asdfa': 0.1, 2: 0.2, 3: 0.3, 4: 0.4}
        dim = Categorical('yolo', categories, shape=2)
        assert 3 not in dim
        assert ('asdfa', 2) in dim

    def test_repr_too_many_cats(self):
        dim = Categorical('yolo', tuple(range(10)), shape=2)
        assert str(dim) == "Categorical(name=yolo, " \
                           "prior={0: 0.10, 1: 0.10, ..., 8: 0.10, 9: 0.10}, " \
                           "shape=(2,))"

    def test_bad_probabilities(self):
        categories = {'asdfa': 0.05, 2: 0.2, 3: 0.3, 4: 0.4}
        with pytest.raises(ValueError):
            Categorical('yolo', categories, shape=2)

    def test_interval_is_banned(self):
        dim = Categorical('yolo', {'asdfa': 0.1, 2: 0.2, 3: 0.3, 4: 0.4}, shape=2)
        with pytest.raises(RuntimeError) as exc:
            dim.interval()
        assert 'not ordered' in str(exc.value)

    @gpu
    def test_that_objects_types_are_ok(self):
        categories = {'asdfa': 0.1, 2: 0.2, 3: 0.3, 'lalala': 0.4}
        dim = Categorical('yolo', categories)
        assert '2' not in dim
        assert 2 in dim
        assert 'asdfa' in dim
        dim = Categorical('yolo', categories, shape=(2,))
        assert ['2', 'asdfa'] not in dim
        assert [2, 'asdfa'] in dim
        for s in dim.sample(20):
            assert s.shape == (2,) and all(type(v) in (str, int) for v in s)


def _three_dim_space(cat_shape=2):
    space = Space()
    probs = (0.1, 0.2, 0.3, 0.4)
    categories = ('asdfa', 2, 3, 4)
    space.register(Categorical('yolo', OrderedDict(zip(categories, probs)), shape=cat_shape))
    space.register(Integer('yolo2', 'uniform', -3, 6))
    space.register(Real('yolo3', 'norm', 0.9))
    return space


class TestSpace(object):

    def test_init(self):
        assert isinstance(Space(), OrderedDict)

    @gpu
    def test_register_and_contain(self):
        space = Space()
        assert 'yolo' not in space
        assert (('asdfa', 2), 0, 3.5) not in space
        space = _three_dim_space()
        assert 'yolo' in space
        assert 'yolo2' in space
        assert 'yolo3' in space
        assert (('asdfa', 2), 0, 3.5) in space
        assert (('asdfa', 2), 7, 3.5) not in space

    def test_bad_contain(self):
        with pytest.raises(TypeError):
            5 in Space()

    @gpu
    def test_sample(self, seed):
        """Ordering = dimension order; one seed drives the whole call (test_space.py:384-417)."""
        space = _three_dim_space(cat_shape=(2, 2))
        dim1, dim2, dim3 = space.values()
        point = space.sample(seed=seed)
        test_point = [(dim1.sample()[0], dim2.sample()[0], dim3.sample()[0])]      # numpy's global state, seeded alike
        assert len(point) == len(test_point) == 1 and len(point[0]) == len(test_point[0]) == 3
        assert np.all(point[0][0] == test_point[0][0])
        assert point[0][1] == test_point[0][1] and point[0][2] == test_point[0][2]
        points = space.sample(2, seed=seed)
        points1, points2, points3 = dim1.sample(2), dim2.sample(2), dim3.sample(2)
        for i in range(2):
            assert np.all(points[i][0] == points1[i])
            assert points[i][1] == points2[i] and points[i][2] == points3[i]
        seed = np.random.RandomState(5)
        point = space.sample(seed=seed)
        assert point[0][0].shape == (2, 2) and point[0] in space
        assert isinstance(point[0][1], np.int64) and isinstance(point[0][2], np.float64)
        points = space.sample(2, seed=seed)
        assert len(points) == 2 and all(len(p) == 3 and p in space for p in points)
        assert all(np.all(x[0] == y[0]) and x[1:] == y[1:]
                   for x, y in zip(space.sample(3, seed=11), space.sample(3, seed=11)))
        np.random.seed(3)
        a = space.sample(2)
        np.random.seed(3)
        b = space.sample(2)
        assert all(np.all(x[0] == y[0]) and x[1:] == y[1:] for x, y in zip(a, b))

    @gpu
    def test_sample_columns_fast_path(self, monkeypatch):
        """The tensor path returns the same values the tuple path converts when both use Philox."""
        from orion_b200 import mt_stream
        monkeypatch.setattr(mt_stream, 'ENABLED', False)
        space = _three_dim_space(cat_shape=(2, 2))
        batch = space.sample_columns(100, seed=4)
        rows = space.rows_from_columns(batch.to_numpy(), 100)
        again = space.sample(100, seed=4)
        assert all(np.all(x[0] == y[0]) and x[1:] == y[1:] for x, y in zip(rows, again))
        assert len(batch.columns) == 6 and batch.columns[0].is_cuda
        assert bool(batch.mask[:100].all().item())
        assert bool(space.contains_columns(batch).all().item())

    def test_interval(self):
        space = _three_dim_space()
        assert space.interval() == [('asdfa', 2, 3, 4), (-3, 3), (-np.inf, np.inf)]

    def test_bad_setitem(self):
        space = Space()
        with pytest.raises(TypeError) as exc:
            space[5] = Integer('yolo', 'uniform', -3, 6)
        assert "string" in str(exc.value)
        with pytest.raises(TypeError) as exc:
            space['ispis'] = 'nope'
        assert "Dimension" in str(exc.value)
        space.register(Integer('yolo', 'uniform', -3, 6))
        with pytest.raises(ValueError) as exc:
            space.register(Real('yolo', 'uniform', 0, 6))
        assert "another name" in str(exc.value)

    def test_getitem(self):
        space = _three_dim_space()
        assert space['yolo'].type == 'categorical'
        assert space[0].type == 'categorical'
        with pytest.raises(KeyError):
            space['asdf']
        with pytest.raises(IndexError):
            space[3]

    def test_repr(self):
        space = Space()
        space.register(Integer('yolo2', 'uniform', -3, 6, shape=(2,)))
        space.register(Real('yolo3', 'norm', 0.9))
        assert str(space) == "Space(["\
                             "Integer(name=yolo2, prior={uniform: (-3, 6), {}}, shape=(2,)),\n" \
                             "       Real(name=yolo3, prior={norm: (0.9,), {}}, shape=())])"


class TestPriorDSL(object):
    """``name~prior(args)`` expressions (reference test_space_builder.py:29-160)."""

    def test_expressions(self):
        from orion_b200.spaces import build_dimension
        dim = build_dimension('lr', 'loguniform(0.001, 1, shape=10)')
        assert str(dim) == "Real(name=lr, prior={reciprocal: (0.001, 1), {}}, shape=(10,))"
        dim = build_dimension('other2', 'random(-5, 2)')
        assert str(dim) == "Real(name=other2, prior={uniform: (-5, 7), {}}, shape=())"
        assert dim.interval() == (-5.0, 2.0)
        assert build_dimension('i', 'uniform(-3, 5, discrete=True)').type == 'integer'
        assert build_dimension('g', 'gaussian(3, 2)')._prior_name == 'norm'
        assert build_dimension('n', 'normal(3, 2)')._prior_name == 'norm'
        c = build_dimension('c', "choices({'a': 0.3, 'b': 0.7})")
        assert c.categories == ('a', 'b') and c._probs == (0.3, 0.7)
        assert build_dimension('c', "choices(['a', 5, 2.5])").categories == ('a', 5, 2.5)
        assert build_dimension('c', "choices('a', 5)").categories == ('a', 5)
        assert build_dimension('p', 'poisson(5)')._prior_name == 'poisson'
        with pytest.raises(TypeError):
            build_dimension('x', 'not an expression')
        with pytest.raises(TypeError):
            build_dimension('x', '__import__("os").system("true")')
        with pytest.raises(ValueError):
            build_dimension('x', 'uniform(__import__("os"), 3)')

    def test_named_spaces(self):
        from orion_b200 import spaces
        s = spaces.mixed64()
        assert len(s) == 64 and [d.type for d in s.values()].count('categorical') == 16
        assert len(spaces.real16()) == 16 and len(spaces.cat32x100()) == 32 and len(spaces.rosen2()) == 2


def test_columns_to_record_batch_schema_and_values():
    """Host half of the bulk trial hand-over (no GPU involved)."""
    import json
    from orion_b200 import format_trials
    from orion_b200.space import Categorical, Integer, Real, Space
    space = Space()
    space.register(Real('lr', 'reciprocal', 1e-5, 1.0))
    space.register(Integer('layers', 'uniform', 1, 8, shape=2))
    space.register(Categorical('act', ['relu', 'tanh', 'gelu']))
    space.register(Categorical('odd', ['asdfa', 2, 3.5]))
    cols = [np.array([0.1, 0.01, 0.5]), np.array([1, 2, 3], dtype=np.int8), np.array([4, 5, 6], dtype=np.int8),
            np.array([2, 0, 1], dtype=np.uint8), np.array([1, 1, 0], dtype=np.uint8)]
    rb = format_trials.columns_to_record_batch(space, cols)
    assert rb.num_rows == 3 and rb.schema.names == ['lr', 'layers[0]', 'layers[1]', 'act', 'odd']
    assert json.loads(rb.schema.metadata[b'orion.types']) == {'lr': 'real', 'layers': 'integer', 'act': 'categorical',
                                                              'odd': 'categorical'}
    assert rb.column(0).to_pylist() == [0.1, 0.01, 0.5]
    assert rb.column(1).to_pylist() == [1, 2, 3] and rb.column(2).to_pylist() == [4, 5, 6]
    assert rb.column(3).to_pylist() == ['gelu', 'relu', 'tanh']
    assert rb.column(4).to_pylist() == ['2', '2', "'asdfa'"]           # mixed-type categories travel as reprs
    back = format_trials.record_batch_to_columns(space, rb)
    assert [c.dtype for c in back] == [np.float64, np.int64, np.int64, np.int32, np.int32]
    assert all(np.array_equal(a, b) for a, b in zip(back, cols))
    with pytest.raises(ValueError):
        format_trials.record_batch_to_columns(space, rb.select([1, 0, 2, 3, 4]))
