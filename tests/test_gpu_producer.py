"""Producer mirror (reference src/orion/core/worker/producer.py:42-67 and
tests/unittests/core/test_producer.py) over an in-memory experiment."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


class ListExperiment(object):
    """What Producer uses of the reference's Experiment, without a database."""

    def __init__(self, space, algorithms, pool_size):
        self.space = space
        self.algorithms = algorithms
        self.pool_size = pool_size
        self.trials = []
        self._completed = []

    def register_trials(self, trials):
        self.trials.extend(trials)

    def fetch_completed_trials(self):
        out, self._completed = self._completed, []
        return out

    def complete(self, trial, objective):
        done = dict(trial, results=[{'name': 'loss', 'type': 'objective', 'value': objective}])
        self._completed.append(done)


def test_produce_registers_pool_size_unique_trials():
    from orion_b200 import spaces
    from orion_b200.primary_algo import PrimaryAlgo
    from orion_b200.producer import Producer
    space = spaces.build_space(['-c~choices([0, 1, 2])', '-i~uniform(0, 4, discrete=True)'])    # 12 distinct points
    exp = ListExperiment(space, PrimaryAlgo(space, {'random': {'seed': 5}}), pool_size=4)
    producer = Producer(exp, max_rounds=64)
    for _ in range(3):
        new = producer.produce()
        assert len(new) == 4
    assert len(exp.trials) == 12
    points = [tuple(p['value'] for p in t['params']) for t in exp.trials]
    assert len(set(points)) == 12                         # the reference would have stored duplicates
    assert [p['name'] for p in exp.trials[0]['params']] == ['c', 'i']
    assert [p['type'] for p in exp.trials[0]['params']] == ['categorical', 'integer']
    assert producer.produce() == []                       # the space is exhausted: nothing new to register


def test_update_observes_results_and_feeds_the_history():
    from orion_b200 import spaces
    from orion_b200.algo.base import BaseAlgorithm
    from orion_b200.producer import Producer

    class Recorder(BaseAlgorithm):
        def __init__(self, space):
            super(Recorder, self).__init__(space)
            self.seen = []

        def suggest(self, num=1):                         # always proposes the same three points
            return [(0.5, 1.5), (2.5, 3.5), (0.5, 1.5)][:num]

        def observe(self, points, results):
            self.seen.extend(zip(points, [r['objective'] for r in results]))

    space = spaces.rosen2()
    algo = Recorder(space)
    exp = ListExperiment(space, algo, pool_size=3)
    producer = Producer(exp, max_rounds=2)
    first = producer.produce()                            # list-of-tuples algorithm: uploaded, deduplicated
    assert [tuple(p['value'] for p in t['params']) for t in first] == [(0.5, 1.5), (2.5, 3.5)]
    exp.complete(first[0], 3.25)
    producer.update()
    assert algo.seen == [((0.5, 1.5), 3.25)]
    assert producer.produce() == []                       # both points are history now
    # a fresh producer learns the history from completed trials alone
    exp2 = ListExperiment(space, Recorder(space), pool_size=3)
    producer2 = Producer(exp2, max_rounds=1)
    exp2.complete(first[1], 1.0)
    producer2.update()
    assert [tuple(p['value'] for p in t['params']) for t in producer2.produce()] == [(0.5, 1.5)]


def test_producer_without_dedup_registers_everything_like_the_reference():
    from orion_b200 import spaces
    from orion_b200.primary_algo import PrimaryAlgo
    from orion_b200.producer import Producer
    space = spaces.build_space(['-c~choices([0, 1])'])
    exp = ListExperiment(space, PrimaryAlgo(space, {'random': {'seed': 1}}), pool_size=8)
    Producer(exp, dedup=False).produce()
    assert len(exp.trials) == 8                           # two categories, eight trials: duplicates registered


def test_batched_score_matches_the_per_point_path_and_uses_one_membership_launch(monkeypatch):
    """§8 f4: Experiment.reserve_trial maps score_handle over every new trial
    (experiment.py:173-178); `score_points` asserts all of them in-space with a
    single library call and returns what per-point `score` returns."""
    from orion_b200 import spaces
    from orion_b200.algo.base import BaseAlgorithm
    from orion_b200.descriptor import CompiledSpace
    from orion_b200.primary_algo import PrimaryAlgo
    from orion_b200.producer import Producer
    from orion_b200 import format_trials

    class Scorer(BaseAlgorithm):
        def __init__(self, space):
            super(Scorer, self).__init__(space)

        def suggest(self, num=1):
            return self.space.sample(num)

        def observe(self, points, results):
            pass

        def score(self, point):                           # prefers small |x|
            return -round(abs(point[0]), 6)

    space = spaces.build_space(['-x~uniform(-5, 10)', '-k~uniform(0, 9, discrete=True)', '-c~choices(["a", "b"])'])
    algo = PrimaryAlgo(space, 'scorer')
    points = space.sample(300, seed=3)
    calls = {'n': 0}
    real = CompiledSpace.contains_host

    def counting(self, *a, **kw):
        calls['n'] += 1
        return real(self, *a, **kw)

    monkeypatch.setattr(CompiledSpace, 'contains_host', counting)
    batched = algo.score_points(points)
    assert calls['n'] == 1                                # one launch for 300 points
    calls['n'] = 0
    single = [algo.score(p) for p in points]
    assert calls['n'] == 300                              # the reference's way: one check per point
    assert batched == single
    judged = algo.judge_points(points[:5], [{}] * 5)
    assert judged == [None] * 5
    # a point outside the space trips the same assertion as the per-point path
    with pytest.raises(AssertionError):
        algo.score_points(points[:3] + [(99.0, 1, 'a')])
    with pytest.raises(AssertionError):
        algo.score((99.0, 1, 'a'))
    assert algo.score_points([]) == []

    # the reservation hook: the best-scored new trial is selected, ties at random
    exp = ListExperiment(space, algo, pool_size=4)
    producer = Producer(exp, dedup=False)
    trials = [format_trials.tuple_to_trial(p, space) for p in points]
    chosen = producer.select_trial(trials)
    best = max(single)
    assert algo.score(format_trials.trial_to_tuple(chosen, space)) == best
    assert producer.select_trial(trials, score_handle=algo.score) is not None
    assert producer.select_trial([]) is None
    with pytest.raises(ValueError):
        producer.select_trial(trials, score_handle=3)
