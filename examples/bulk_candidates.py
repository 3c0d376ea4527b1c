#!/usr/bin/env python
"""The tensor fast path: ten million candidates of a mixed space, transformed for
an algorithm that needs real inputs, reversed, and deduplicated against a trial
history -- all device-resident (stream pipeline), then the survivors gathered
and copied to the host once.

    python examples/bulk_candidates.py
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from orion_b200 import spaces  # noqa: E402
from orion_b200.history import TrialHistory  # noqa: E402
from orion_b200.pipeline import CandidatePipeline  # noqa: E402
from orion_b200.transformer import build_required_space  # noqa: E402


def main():
    n = 10_000_000
    space = spaces.mixed64()
    tspace = build_required_space('real', space)
    compiled = tspace.compiled()
    history = TrialHistory(compiled, capacity=8 * n)
    pipe = CandidatePipeline(compiled, history, n, fingerprint_in='suggest')      # one kernel per batch + dedup
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    slots = []
    for step in range(4):
        slot = pipe.submit(seed=2024, base_index=(step % 3) * n)      # the 4th batch repeats the 1st: all duplicates
        pipe.drain()
        slots.append(int(slot.batch.keep[:n].sum().item()))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("kept per batch:", slots, "| %.2e candidates/s" % (4 * n / dt))
    slot = pipe.submit(seed=7, base_index=0)
    pipe.drain()
    survivors = slot.back.compact(slot.batch.keep)                      # dense device block of the kept rows
    columns = survivors.to_numpy()                                      # one host copy
    print("survivors on the host:", survivors.n, "rows x", len(columns), "columns;",
          "transformed width:", compiled.n_tcols)


if __name__ == '__main__':
    main()
