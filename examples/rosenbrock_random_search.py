#!/usr/bin/env python
"""BASELINE.json config 1 in process: random search on the 2-dim Rosenbrock
function through the reference-shaped API (PrimaryAlgo -> Random -> Space.sample),
with trials registered through the Producer mirror.  Needs a B200 and libob2.so.

    python examples/rosenbrock_random_search.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from orion_b200 import spaces  # noqa: E402
from orion_b200.primary_algo import PrimaryAlgo  # noqa: E402
from orion_b200.producer import Producer  # noqa: E402


def rosenbrock(x, y):
    return (1 - x) ** 2 + 100 * (y - x ** 2) ** 2


class Experiment(object):
    """The slice of the reference's Experiment that Producer talks to."""

    def __init__(self, space, algorithms, pool_size):
        self.space, self.algorithms, self.pool_size = space, algorithms, pool_size
        self.trials, self._done = [], []

    def register_trials(self, trials):
        self.trials.extend(trials)

    def fetch_completed_trials(self):
        done, self._done = self._done, []
        return done

    def evaluate(self, trial):
        x, y = (p['value'] for p in trial['params'])
        self._done.append(dict(trial, results=[{'name': 'rosenbrock', 'type': 'objective', 'value': rosenbrock(x, y)}]))
        return self._done[-1]['results'][0]['value']


def main():
    space = spaces.build_space(['-x~uniform(-5, 10)', '-y~uniform(-5, 10)'])
    experiment = Experiment(space, PrimaryAlgo(space, {'random': {'seed': 1}}), pool_size=10)
    producer = Producer(experiment)
    best = float('inf')
    for _ in range(100):
        for trial in producer.produce():
            best = min(best, experiment.evaluate(trial))
        producer.update()
    print("%d trials, best objective %.4f" % (len(experiment.trials), best))


if __name__ == '__main__':
    main()
