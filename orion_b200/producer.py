"""Producer: suggest -> dedup -> register, observe completed trials.

Mirror of ``src/orion/core/worker/producer.py:18-67`` for the callers either
side of the hot path.  The reference's ``produce`` registers every suggestion
unconditionally (``tests/unittests/core/test_producer.py:69-88`` shows two
identical points both stored); here suggestions are deduplicated against the
trial history on the device before they are turned into trial records, and
``update`` feeds completed trials back into that history -- the way upstream
Orion's producer relies on the database's unique trial hash.

The ``experiment`` argument is duck-typed on what the reference's Producer uses
of it: ``pool_size``, ``space``, ``algorithms``, ``register_trials(trials)`` and
``fetch_completed_trials()`` (``src/orion/core/worker/experiment.py:216-247``).
Trials are the plain records of :mod:`orion_b200.format_trials`.
"""
import logging
import random

from . import format_trials
from .history import TrialHistory

log = logging.getLogger(__name__)


def get_trial_results(trial):
    """``{'objective', 'constraint', 'gradient'}`` of a completed trial record
    (reference ``format_trials.get_trial_results``, format_trials.py:45-55)."""
    results = trial.get('results', [])
    objective = [r['value'] for r in results if r.get('type') == 'objective']
    gradient = [r['value'] for r in results if r.get('type') == 'gradient']
    return {
        'objective': objective[0] if objective else None,
        'constraint': [r['value'] for r in results if r.get('type') == 'constraint'],
        'gradient': tuple(gradient[0]) if gradient else None,
    }


class Producer(object):
    """Produce new, never-seen trials; keep the algorithm and the history current."""

    def __init__(self, experiment, dedup=True, history_capacity=1 << 20, max_rounds=16):
        log.debug("Creating Producer object.")
        self.experiment = experiment
        self.num_new_trials = experiment.pool_size
        self.space = experiment.space
        if self.space is None:
            raise RuntimeError("Experiment object provided to Producer has not yet completed initialization.")
        self.algorithm = experiment.algorithms
        self.dedup = dedup
        self.max_rounds = max_rounds
        self.history = TrialHistory(self.space, capacity=history_capacity, auto_grow=True) if dedup else None

    def _suggest_batch(self, num):
        """Device batch of ``num`` suggestions: straight from the algorithm's
        tensor path when it has one, else its list of tuples uploaded."""
        if hasattr(self.algorithm, 'suggest_columns'):
            return self.algorithm.suggest_columns(num)
        points = self.algorithm.suggest(num)
        return format_trials.trials_to_batch(self.space, [format_trials.tuple_to_trial(p, self.space) for p in points])

    def produce(self):
        """Create and register ``pool_size`` new trials (fewer only if the
        algorithm keeps proposing known points for ``max_rounds`` rounds)."""
        wanted = self.num_new_trials
        new_trials = []
        for _ in range(self.max_rounds):
            if len(new_trials) >= wanted:
                break
            batch = self._suggest_batch(wanted - len(new_trials))
            if self.dedup:
                self.history.dedup(batch)                  # sets batch.keep; survivors join the history
            new_trials.extend(format_trials.batch_to_trials(self.space, batch))
        new_trials = new_trials[:wanted]
        log.debug("### Register to database: %s", new_trials)
        self.experiment.register_trials(new_trials)
        return new_trials

    def update(self):
        """Pull newly completed trials: record them in the history and let the
        algorithm observe their results."""
        completed = self.experiment.fetch_completed_trials()
        if not completed:
            return
        points = [format_trials.trial_to_tuple(t, self.space) for t in completed]
        results = [get_trial_results(t) for t in completed]
        if self.dedup:
            self.history.observe(format_trials.trials_to_batch(self.space, completed))
        self.algorithm.observe(points, results)

    def select_trial(self, new_trials, score_handle=None):
        """The scoring half of ``Experiment.reserve_trial``
        (src/orion/core/worker/experiment.py:170-183): among the *new* trial records pick one
        with the highest score, at random among ties.  ``score_handle`` defaults to the
        algorithm's batched ``score_points`` -- every candidate's ``point in space`` assertion
        in one kernel launch -- and may be any callable taking one point, as in the
        reference (``workon`` passes ``producer.algorithm.score``, worker/__init__.py:29).
        The database half (find / find_one_and_update) stays with the experiment object."""
        if score_handle is not None and not callable(score_handle):
            raise ValueError("Argument `score_handle` must be callable with a `Trial`.")
        new_trials = list(new_trials)
        if not new_trials:
            return None
        points = [format_trials.trial_to_tuple(t, self.space) for t in new_trials]
        if score_handle is None and hasattr(self.algorithm, 'score_points'):
            scores = list(self.algorithm.score_points(points))
        else:
            handle = score_handle if score_handle is not None else self.algorithm.score
            scores = [handle(p) for p in points]
        best = max(scores)
        return random.choice([t for s, t in zip(scores, new_trials) if s == best])
