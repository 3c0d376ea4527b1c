"""Random search on the device.

Mirror of ``src/orion/algo/random.py:15-41`` (``suggest(num)`` =
``space.sample(num)``), plus:

* ``suggest_columns(num)``: the tensor-returning fast path (a ColumnBatch whose
  ``mask`` is the fused ``point in space`` and, with a history attached,
  ``keep`` the dedup verdict);
* optional dedup against the trial history: with
  ``dedup=True`` the list path tops up until ``num`` unique points.
"""
import numpy

from .._lib import HOST_SAMPLE_MAX
from ..space import _raise_if_failed, philox_key
from .base import BaseAlgorithm


class Random(BaseAlgorithm):
    """Sample candidates independently from the space's priors."""

    def __init__(self, space, seed=None, dedup=False, history_capacity=1 << 20):
        super(Random, self).__init__(space, seed=seed, dedup=dedup, history_capacity=history_capacity)
        self._rng = numpy.random.RandomState(seed) if seed is not None else None
        self._history = None
        self._next_index = 0

    @property
    def history(self):
        """Lazily created :class:`orion_b200.history.TrialHistory`."""
        if self._history is None:
            from ..history import TrialHistory
            # rows are remembered as they are handed out in the ORIGINAL space (what a trial stores)
            self._history = TrialHistory(getattr(self.space, 'original', self.space), capacity=self.history_capacity)
        return self._history

    @property
    def _transformed(self):
        """Is the algorithm's space a TransformedSpace (the way PrimaryAlgo is meant to wrap it)?"""
        return hasattr(self.space, 'original')

    def suggest_columns(self, num=1, check=True):
        """Device-resident suggestion: ColumnBatch with ``mask`` (and ``keep``
        when dedup is on; kept rows join the history)."""
        if self._transformed:
            # one kernel: sample -> transform -> reverse -> fingerprint.  The algorithm's points are the
            # TRANSFORMED batch; dedup is on the rows as handed out (reverse of it), kept in ``.original``.
            batch, tbatch, back = self.space.suggest_columns(num, self._rng, check=check)
            tbatch.mask, tbatch.fp, tbatch.original = batch.mask, back.fp, back
            if self.dedup:
                tbatch.keep = self.history.dedup_fingerprints(back.fp, num)
            return tbatch
        batch = self.space.sample_columns(num, self._rng, with_fingerprint=bool(self.dedup), check=check)
        if self.dedup:
            self.history.dedup(batch)
        return batch

    def suggest(self, num=1):
        """``num`` points as a list of tuples.  Without dedup this is exactly
        ``space.sample(num)`` (seed None -> numpy's global RNG)."""
        if not self.dedup:
            return self.space.sample(num, self._rng)
        points = []
        for _ in range(128):                            # bounded top-up (tiny discrete spaces need many rounds)
            need = num - len(points)
            if need <= 0:
                break
            if self._transformed:
                tbatch = self.suggest_columns(need)
                keep = tbatch.keep[:need].cpu().numpy().astype(bool)
                rows = self.space._rows_from_tcols(tbatch.to_numpy(), need)
            elif need <= HOST_SAMPLE_MAX:               # draw + dedup + copy + sync in one library call
                compiled = self.space.compiled()
                cols, failed, keep = compiled.sample_host(need, philox_key(self._rng), history=self.history)
                _raise_if_failed(failed, compiled, list(self.space.values()))
                rows = self.space.rows_from_columns(cols, need)
            else:
                batch = self.suggest_columns(need)
                keep = batch.keep[:batch.n].cpu().numpy().astype(bool)
                rows = self.space.rows_from_columns(batch.to_numpy(), batch.n)
            points.extend(r for r, k in zip(rows, keep) if k)
        return points[:num]

    def observe(self, points, results):
        """Random search learns nothing from results; with dedup on, observed
        points are recorded so they are never suggested again."""
        if self.dedup and points:
            if self._transformed:
                points = self.space.reverse_points(points)
            self.history.observe_points(points)
