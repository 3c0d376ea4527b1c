#!/usr/bin/env python
"""Measurement of the widened rows on one B200, mixed64, with the
unmodified reference's own per-point code (oracle/_ref) timed beside each where it has some:

  f1  survivors -> trial records:   batch_to_record_batch (device gather, one copy, arrow columns) and
                                    batch_to_trials (Python records); reference: tuple_to_trial per point
                                    (src/orion/core/utils/format_trials.py:28-42)
  f2  stored trials -> history:     record_batch_to_columns -> ColumnBatch -> TrialHistory.observe, and
                                    trials_to_batch -> observe; reference: trial_to_tuple per trial (:16-25)
  f4  scoring a list of points:     PrimaryAlgo.score_points (one membership launch); reference:
                                    PrimaryAlgo.score per point (src/orion/core/worker/primary_algo.py:75-84)

    python benchmarks/widened.py > gpurun_out/widened.json
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from orion_b200 import spaces  # noqa: E402
from orion_b200.columns import ColumnBatch  # noqa: E402
from orion_b200.format_trials import (batch_to_record_batch, batch_to_trials, record_batch_to_columns,  # noqa: E402
                                      trials_to_batch)
from orion_b200.history import TrialHistory  # noqa: E402
from orion_b200.primary_algo import PrimaryAlgo  # noqa: E402


def wall(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = float('inf')
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best


def main():
    out = {}
    space = spaces.mixed64()
    compiled = space.compiled()
    n_bulk, n_records = 1_000_000, 20_000

    # ---- f1: survivors -> records
    hist = TrialHistory(space, capacity=8 * n_bulk)
    batch = compiled.sample(n_bulk, 11, with_fingerprint=True)
    hist.observe(compiled.sample(n_bulk // 10, 11, with_fingerprint=True))        # the first tenth is already known
    hist.dedup(batch)
    kept = int(batch.keep[:n_bulk].sum().item())
    dt = wall(lambda: batch_to_record_batch(space, batch))
    rb = batch_to_record_batch(space, batch)
    out['f1_batch_to_record_batch'] = {'rows': kept, 'seconds': dt, 'rows_per_s': kept / dt, 'bytes': rb.nbytes,
                                       'note': 'ob2_compact_wire on the device, one device->host copy, arrow columns'}
    small = compiled.sample(n_records, 12, with_fingerprint=True)
    dt = wall(lambda: batch_to_trials(space, small), 2)
    out['f1_batch_to_trials'] = {'rows': n_records, 'seconds': dt, 'rows_per_s': n_records / dt,
                                 'note': "one {'name','type','value'} dict per parameter, as Trial.to_dict stores them"}

    # ---- f2: stored trials -> history
    def ingest_bulk():
        h = TrialHistory(space, capacity=4 * kept)
        h.observe(ColumnBatch.from_numpy(compiled, record_batch_to_columns(space, rb)))
        return h
    dt = wall(ingest_bulk)
    assert len(ingest_bulk()) == kept
    out['f2_record_batch_to_history'] = {'rows': kept, 'seconds': dt, 'rows_per_s': kept / dt,
                                         'note': 'arrow columns -> device -> fingerprint -> hash set insert'}
    trials = batch_to_trials(space, small)

    def ingest_records():
        h = TrialHistory(space, capacity=4 * n_records)
        h.observe(trials_to_batch(space, trials))
        return h
    dt = wall(ingest_records, 2)
    out['f2_trials_to_history'] = {'rows': n_records, 'seconds': dt, 'rows_per_s': n_records / dt}

    # ---- f4: scoring a list
    algo = PrimaryAlgo(space, 'random')
    points = space.sample(n_records, seed=5)
    dt = wall(lambda: algo.score_points(points), 2)
    out['f4_score_points'] = {'points': n_records, 'seconds': dt, 'points_per_s': n_records / dt,
                              'note': 'one ob2_contains launch for the whole list'}
    dt = wall(lambda: [algo.score(p) for p in points[:500]], 2)
    out['f4_score_one_by_one'] = {'points': 500, 'seconds': dt, 'points_per_s': 500 / dt,
                                  'note': 'the same through score(point): one ob2_contains_host call each'}

    # ---- the reference's per-point code on one host core
    try:
        from oracle import ref_harness
        ref_space = ref_harness.build_space(ref_harness.MIXED64_ARGS)
        from orion.core.utils.format_trials import trial_to_tuple as ref_trial_to_tuple
        from orion.core.utils.format_trials import tuple_to_trial as ref_tuple_to_trial
        ref_points = ref_space.sample(300, seed=5)
        t0 = time.perf_counter()
        ref_trials = [ref_tuple_to_trial(p, ref_space) for p in ref_points]
        dt = time.perf_counter() - t0
        out['f1_reference_tuple_to_trial'] = {'rows': 300, 'seconds': dt, 'rows_per_s': 300 / dt}
        t0 = time.perf_counter()
        back = [ref_trial_to_tuple(t, ref_space) for t in ref_trials]
        dt = time.perf_counter() - t0
        assert len(back) == 300
        out['f2_reference_trial_to_tuple'] = {'rows': 300, 'seconds': dt, 'rows_per_s': 300 / dt}
        ref_algo = ref_harness.primary_algo(ref_space, 'random')
        t0 = time.perf_counter()
        for p in ref_points:
            ref_algo.score(p)
        dt = time.perf_counter() - t0
        out['f4_reference_score'] = {'points': 300, 'seconds': dt, 'points_per_s': 300 / dt}
    except Exception as exc:  # noqa: BLE001 -- the reference is optional here
        out['reference'] = {'unavailable': '%s: %s' % (type(exc).__name__, exc)}
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
