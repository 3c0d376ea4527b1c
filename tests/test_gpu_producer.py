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
