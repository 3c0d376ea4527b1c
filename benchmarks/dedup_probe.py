"""Experiment: how does ob2_set_dedup's time depend on the table's footprint and on
the locality of the probes?  (run under gpurun)

For each table size: fill to ~25 % with random fingerprints, then dedup batches of
1.25e7 fresh random fingerprints (a) in random order, (b) sorted by slot position
(upper bound of what partitioning the probes by table region could buy).
"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from orion_b200 import spaces
from orion_b200.history import TrialHistory

dev = torch.device('cuda', 0)
compiled = spaces.rosen2().compiled()
n = 12_500_000
g = torch.Generator(device=dev)
g.manual_seed(1)


def rand_fp(m):
    fp = torch.randint(-(1 << 62), 1 << 62, (m, 2), generator=g, device=dev, dtype=torch.int64)
    fp[:, 0] |= 1          # lo != 0
    return fp


def timed(fn, reps=1):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for log2 in (28, 29, 30, 31):
    slots = 1 << log2
    hist = TrialHistory(compiled, n_slots=slots)
    fill = min(slots // 8, 200_000_000)
    done = 0
    while done < fill:
        m = min(n, fill - done)
        hist.insert_fingerprints(rand_fp(m))
        done += m
    keep = torch.empty(n, dtype=torch.uint8, device=dev)
    res = []
    for mode in ('random', 'sorted', 'part256', 'part4096'):
        ts = []
        for rep in range(2):
            fp = rand_fp(n)
            if mode != 'random':
                pos = fp[:, 0] & (slots - 1)
                if mode == 'sorted':
                    key = pos
                else:
                    parts = 256 if mode == 'part256' else 4096
                    key = pos >> (log2 - parts.bit_length() + 1)
                fp = fp[torch.argsort(key)].contiguous()
            ts.append(timed(lambda: hist.dedup_fingerprints(fp, n, keep=keep)))
        res.append('%s %.3f' % (mode, min(ts)))
    print('table 2^%d slots (%.1f GB), filled %.0f%%: dedup of %d fresh: %s ms' % (
        log2, slots * 32 / 1e9, 100.0 * len(hist) / slots, n, '  '.join(res)), flush=True)
    del hist, keep
    torch.cuda.empty_cache()
