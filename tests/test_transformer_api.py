"""TransformedSpace / build_required_space.

The reference's own ``TestPrimaryAlgoTransforms`` is an empty class
(tests/unittests/core/test_primary_algo.py:107-110): these are builder-authored
known-answer tests written from the worked examples of Appendix B.1.
"""
from collections import OrderedDict

import numpy as np
import pytest

from orion_b200.space import Categorical, Integer, Real, Space
from orion_b200.transformer import (Compose, Enumerate, Identity, OneHotEncode, Quantize, Reverse,
                                    TransformedSpace, build_required_space)

gpu = pytest.mark.gpu


@pytest.fixture()
def space():
    space = Space()
    space.register(Categorical('yolo', OrderedDict(zip(('asdfa', 2, 3, 4), (0.1, 0.2, 0.3, 0.4))), shape=2))
    space.register(Integer('yolo2', 'uniform', -3, 6))
    space.register(Real('yolo3', 'norm', 0.9))
    space.register(Categorical('yolo4', ('a', 'b')))
    return space


class TestChainDescription(object):
    """Types, shapes and intervals of the chain -- no device needed."""

    def test_transformer_types(self):
        assert (Quantize().domain_type, Quantize().target_type) == ('real', 'integer')
        r = Reverse(Quantize())
        assert (r.domain_type, r.target_type) == ('integer', 'real')
        e = Enumerate(('a', 'b', 'c'))
        assert (e.domain_type, e.target_type) == ('categorical', 'integer') and e.interval() == (0, 3)
        o = OneHotEncode(4)
        assert (o.domain_type, o.target_type) == ('integer', 'real')
        assert o.infer_target_shape(()) == (4,) and o.infer_target_shape((2,)) == (2, 4)
        assert OneHotEncode(2).infer_target_shape((3,)) == (3,)
        c = Compose([e, OneHotEncode(3)])
        assert (c.domain_type, c.target_type) == ('categorical', 'real')
        assert c.infer_target_shape((2,)) == (2, 3)
        assert Identity('real').target_type == 'real'

    def test_build_required_space_table(self, space):
        real = build_required_space('real', space)
        assert isinstance(real, TransformedSpace)
        assert [d.type for d in real.values()] == ['real', 'real', 'real', 'real']
        assert [d.shape for d in real.values()] == [(2, 4), (), (), ()]
        assert [d.name for d in real.values()] == ['yolo', 'yolo2', 'yolo3', 'yolo4']
        integer = build_required_space('integer', space)
        assert [d.type for d in integer.values()] == ['integer', 'integer', 'integer', 'integer']
        assert [d.shape for d in integer.values()] == [(2,), (), (), ()]
        numerical = build_required_space('numerical', space)
        assert [d.type for d in numerical.values()] == ['real', 'integer', 'real', 'real']
        none = build_required_space(None, space)
        assert [d.type for d in none.values()] == ['categorical', 'integer', 'real', 'categorical']
        with pytest.raises(TypeError):
            build_required_space('fasdfasf', space)

    def test_intervals(self, space):
        real = build_required_space('real', space)
        assert real['yolo2'].interval() == (-3, 3)
        assert real['yolo4'].interval() == (0.0, 1.0)
        assert real['yolo'].interval() == ((0.0,) * 4, (1.0,) * 4)
        assert build_required_space('integer', space)['yolo'].interval() == (0, 4)


@gpu
class TestOnDevice(object):

    def test_worked_examples(self):
        """B.1: k=4, i=2 -> [0,0,1,0]; reverse([0.1,0.7,0.7,0.0]) -> 1."""
        s = Space()
        s.register(Categorical('c', (10, 11, 12, 13)))
        t = build_required_space('real', s)
        assert np.array_equal(t.transform((12,))[0], [0., 0., 1., 0.])
        assert t.reverse((np.array([0.1, 0.7, 0.7, 0.0]),)) == (11,)
        s = Space()
        s.register(Real('r', 'uniform', -10, 20))
        q = build_required_space('integer', s)
        assert [q.transform((v,))[0] for v in (0.5, 1.5, 2.5, -0.5, 2.4)] == [0, 2, 2, 0, 2]   # half to even
        assert isinstance(q.reverse((3,))[0], np.float64) and q.reverse((3,))[0] == 3.0
        qf = build_required_space('integer', s, quantize='floor')
        assert [qf.transform((v,))[0] for v in (0.5, 1.5, -0.5)] == [0, 1, -1]

    def test_point_round_trip(self, space):
        for requirement in ('real', 'numerical', None):
            tspace = build_required_space(requirement, space)
            for point in space.sample(20, seed=1):
                tpoint = tspace.transform(point)
                assert tpoint in tspace
                back = tspace.reverse(tpoint)
                assert np.all(back[0] == point[0]) and back[1:] == point[1:]

    def test_transformed_point_shapes_and_types(self, space):
        tspace = build_required_space('real', space)
        point = (np.array(['asdfa', 3], dtype=object), 2, 0.75, 'b')
        t = tspace.transform(point)
        assert t[0].shape == (2, 4) and np.array_equal(t[0], [[1, 0, 0, 0], [0, 0, 1, 0]])
        assert isinstance(t[1], np.float64) and t[1] == 2.0
        assert t[2] == 0.75 and t[3] == 1.0
        ispace = build_required_space('integer', space)
        t = ispace.transform(point)
        assert np.array_equal(t[0], [0, 2]) and t[1] == 2 and t[2] == 1 and t[3] == 1
        with pytest.raises(AssertionError):
            tspace.transform((np.array(['nope', 3], dtype=object), 2, 0.75, 'b'))

    def test_sample_and_contains(self, space):
        tspace = build_required_space('real', space)
        points = tspace.sample(50, seed=2)
        assert len(points) == 50 and all(p in tspace for p in points)
        originals = space.sample(50, seed=2)
        again = tspace.transform_points(originals)
        for a, b in zip(points, again):           # sample == transform(original.sample) (same seed)
            assert np.array_equal(a[0], b[0]) and a[1:] == b[1:]
        outside = (np.zeros((2, 4)), 7.0, 0.5, 0.0)   # reverse -> yolo2 == 7: outside [-3, 3)
        assert outside not in tspace
        assert 'yolo' in tspace and 'nope' not in tspace

    def test_dimension_level_calls(self, space):
        tspace = build_required_space('real', space)
        tdim = tspace['yolo4']
        assert tdim.transform('b') == 1.0 and tdim.reverse(0.2) == 'a'
        assert 0.9 in tdim
        assert all(v in (0.0, 1.0) for v in tdim.sample(10, seed=0))
