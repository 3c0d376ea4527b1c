"""BaseAlgorithm / OptimizationAlgorithm factory / PrimaryAlgo / Random behave
like the reference's (tests/unittests/algo/test_base.py,
tests/unittests/core/test_primary_algo.py, with the DumbAlgo stub of
tests/conftest.py:13-70)."""
from collections import OrderedDict

import numpy as np
import pytest

from orion_b200.algo.base import BaseAlgorithm, OptimizationAlgorithm
from orion_b200.primary_algo import PrimaryAlgo
from orion_b200.space import Categorical, Integer, Real, Space

gpu = pytest.mark.gpu


class DumbAlgo(BaseAlgorithm):
    """Stub algorithm: returns a fixed value, logs what it is given."""

    def __init__(self, space, value=5, scoring=0, judgement=None, suspend=False, done=False, **nested_algo):
        self._times_called_suspend = 0
        self._times_called_is_done = 0
        self._num = None
        self._points = None
        self._results = None
        self._score_point = None
        self._judge_point = None
        self._measurements = None
        super(DumbAlgo, self).__init__(space, value=value, scoring=scoring, judgement=judgement,
                                       suspend=suspend, done=done, **nested_algo)

    def suggest(self, num=1):
        self._num = num
        return [self.value] * num

    def observe(self, points, results):
        self._points = points
        self._results = results

    def score(self, point):
        self._score_point = point
        return self.scoring

    def judge(self, point, measurements):
        self._judge_point = point
        self._measurements = measurements
        return self.judgement

    @property
    def should_suspend(self):
        self._times_called_suspend += 1
        return self.suspend

    @property
    def is_done(self):
        self._times_called_is_done += 1
        return self.done


@pytest.fixture()
def space():
    """One of each dimension kind (tests/unittests/core/conftest.py:60-71)."""
    space = Space()
    categories = OrderedDict(zip(('asdfa', 2, 3, 4), (0.1, 0.2, 0.3, 0.4)))
    space.register(Categorical('yolo', categories, shape=2))
    space.register(Integer('yolo2', 'uniform', -3, 6))
    space.register(Real('yolo3', 'alpha', 0.9))
    return space


@pytest.fixture()
def fixed_suggestion():
    return (('asdfa', 2), 0, 3.5)


@pytest.fixture()
def palgo(space, fixed_suggestion):
    algo_config = {'DumbAlgo': {'value': fixed_suggestion,
                                'subone': {'DumbAlgo': dict(value=6, scoring=5)}}}
    return PrimaryAlgo(space, algo_config)


class TestBase(object):
    """tests/unittests/algo/test_base.py"""

    def test_init(self):
        nested = DumbAlgo(8, value=1)
        algo = DumbAlgo(8, value=1, naedw=nested)
        assert algo.space == 8 and algo.value == 1 and algo.naedw is nested
        algo = DumbAlgo(8, value=1, naedw={'DumbAlgo': dict(value=9)})
        assert isinstance(algo.naedw, DumbAlgo) and algo.naedw.value == 9 and algo.naedw.space == 8
        algo = DumbAlgo(8, value=1, naedw='dumbalgo')
        assert isinstance(algo.naedw, DumbAlgo)

    def test_configuration(self):
        algo = DumbAlgo(8, value=1, naedw={'DumbAlgo': dict(value=9, scoring=5)})
        assert algo.configuration == {'dumbalgo': {
            'value': 1, 'scoring': 0, 'judgement': None, 'suspend': False, 'done': False,
            'naedw': {'dumbalgo': {'value': 9, 'scoring': 5, 'judgement': None, 'suspend': False, 'done': False}}}}

    def test_space_setter(self):
        algo = DumbAlgo(8, value=1, naedw={'DumbAlgo': dict(value=9)})
        algo.space = 'etsh'
        assert algo.space == 'etsh' and algo.naedw.space == 'etsh'

    def test_factory(self):
        assert isinstance(OptimizationAlgorithm('dumbalgo', 3), DumbAlgo)
        assert 'random' in OptimizationAlgorithm.typenames or OptimizationAlgorithm('random', Space()) is not None
        with pytest.raises(NotImplementedError) as exc:
            OptimizationAlgorithm('nope', 3)
        assert "Could not find implementation of BaseAlgorithm, type = 'nope'" in str(exc.value)


@gpu
class TestPrimaryAlgoWraps(object):
    """tests/unittests/core/test_primary_algo.py:26-104"""

    def test_init_and_configuration(self, palgo, fixed_suggestion):
        assert isinstance(palgo.algorithm, DumbAlgo)
        assert palgo.configuration == {'dumbalgo': {
            'value': fixed_suggestion, 'scoring': 0, 'judgement': None, 'suspend': False, 'done': False,
            'subone': {'dumbalgo': {'value': 6, 'scoring': 5, 'judgement': None, 'suspend': False, 'done': False}}}}

    def test_space_can_only_retrieved(self, palgo, space):
        assert palgo.space == space
        with pytest.raises(AttributeError):
            palgo.space = 5

    def test_suggest(self, palgo, fixed_suggestion):
        assert palgo.suggest() == [fixed_suggestion]
        assert palgo.suggest(4) == [fixed_suggestion] * 4
        palgo.algorithm.value = (5,)
        with pytest.raises(AssertionError):
            palgo.suggest()

    def test_observe(self, palgo, fixed_suggestion):
        palgo.observe([fixed_suggestion], [5])
        assert palgo.algorithm._points == [fixed_suggestion]
        assert palgo.algorithm._results == [5]
        with pytest.raises(AssertionError):
            palgo.observe([fixed_suggestion], [5, 8])
        with pytest.raises(AssertionError):
            palgo.observe([(5,)], [5])

    def test_isdone(self, palgo):
        palgo.algorithm.done = 10
        assert palgo.is_done == 10
        assert palgo.algorithm._times_called_is_done == 1

    def test_shouldsuspend(self, palgo):
        palgo.algorithm.suspend = 55
        assert palgo.should_suspend == 55
        assert palgo.algorithm._times_called_suspend == 1

    def test_score(self, palgo, fixed_suggestion):
        palgo.algorithm.scoring = 60
        assert palgo.score(fixed_suggestion) == 60
        assert palgo.algorithm._score_point == fixed_suggestion
        with pytest.raises(AssertionError):
            palgo.score((5,))

    def test_judge(self, palgo, fixed_suggestion):
        palgo.algorithm.judgement = 'naedw'
        assert palgo.judge(fixed_suggestion, 8) == 'naedw'
        assert palgo.algorithm._judge_point == fixed_suggestion
        assert palgo.algorithm._measurements == 8
        with pytest.raises(AssertionError):
            palgo.judge((5,), 8)


@gpu
class TestRandom(object):
    """src/orion/algo/random.py:24-33 through PrimaryAlgo (BASELINE config 1:
    random search on a 2-dim Rosenbrock, in process)."""

    def test_suggest_is_space_sample(self):
        from orion_b200 import spaces
        space = spaces.mixed64()
        algo = PrimaryAlgo(space, 'random')
        np.random.seed(1)
        a = algo.suggest(5)
        np.random.seed(1)
        b = space.sample(5)
        assert a == b and len(a) == 5 and all(p in space for p in a)

    def test_rosenbrock_random_search(self):
        from orion_b200 import spaces
        space = spaces.rosen2()
        algo = PrimaryAlgo(space, {'random': {'seed': 3}})
        best = np.inf
        for _ in range(100):
            points = algo.suggest(10)
            results = [{'objective': (1 - x) ** 2 + 100 * (y - x ** 2) ** 2} for x, y in points]
            algo.observe(points, results)
            best = min(best, min(r['objective'] for r in results))
        assert best < 5.0

    def test_suggest_columns_and_dedup(self):
        from orion_b200 import spaces
        space = spaces.build_space(['-c~choices([0, 1, 2])', '-i~uniform(0, 4, discrete=True)'])   # 12 distinct points
        algo = PrimaryAlgo(space, {'random': {'seed': 0, 'dedup': True, 'history_capacity': 4096}})
        points = algo.suggest(12)            # tops up until unique
        assert len(points) == len(set(points)) == 12
        batch = algo.suggest_columns(1000)   # everything is a duplicate now
        assert not bool(batch.keep[:1000].any().item())
        algo2 = PrimaryAlgo(space, {'random': {'seed': 0, 'dedup': True}})
        algo2.observe(points[:6], [{'objective': 0.0}] * 6)
        more = algo2.suggest(6)
        assert len(more) == 6 and not set(more) & set(points[:6])
