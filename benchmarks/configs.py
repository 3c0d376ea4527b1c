#!/usr/bin/env python
"""Side measurements for the BASELINE.json configs that are not the bench line
(config 1: small-n latency through the list-of-tuples API; config 2: real16
sampling; config 4: cat32x100 one-hot transform + reverse, chunked).  Prints one
JSON object; numbers go to DESIGN.md, not to the driver."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from orion_b200 import spaces  # noqa: E402
from orion_b200.primary_algo import PrimaryAlgo  # noqa: E402
from orion_b200.transformer import build_required_space  # noqa: E402


def timed(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    out = {}
    # config 1: random search on 2-dim Rosenbrock, suggest(10) x 100, in process
    # (seed None: numpy's global state, as the reference's Random draws -- the kernel is fed that stream, so both
    # loops below see the same 1000 points after the same numpy.random.seed)
    import numpy
    algo = PrimaryAlgo(spaces.rosen2(), 'random')
    algo.suggest(10)
    numpy.random.seed(1)
    t0 = time.perf_counter()
    best = float('inf')
    mine = []
    for _ in range(100):
        pts = algo.suggest(10)
        mine.extend(pts)
        best = min(best, min((1 - x) ** 2 + 100 * (y - x ** 2) ** 2 for x, y in pts))
    dt = time.perf_counter() - t0
    out['config1_rosen2_suggest10_x100'] = {'seconds': dt, 'candidates_per_s': 1000 / dt, 'best_objective': best,
                                            'note': "list-of-tuples API on the reference's own MT19937 stream: numbers in, "
                                                    "launch, D2H, Python tuples per call"}
    # the same loop through the UNMODIFIED reference (oracle/_ref), for the record: it is config 1's own CPU path
    try:
        from oracle import ref_harness
        ref_space = ref_harness.build_space(['-x~uniform(-5, 10)', '-y~uniform(-5, 10)'])
        ref_algo = ref_harness.primary_algo(ref_space, 'random')
        ref_algo.suggest(10)
        numpy.random.seed(1)
        t0 = time.perf_counter()
        ref_best = float('inf')
        theirs = []
        for _ in range(100):
            rp = ref_algo.suggest(10)
            theirs.extend(rp)
            ref_best = min(ref_best, min((1 - x) ** 2 + 100 * (y - x ** 2) ** 2 for x, y in rp))
        rdt = time.perf_counter() - t0
        out['config1_same_1000_points_as_the_reference'] = bool(mine == theirs)
        out['config1_reference_cpu_path'] = {'seconds': rdt, 'candidates_per_s': 1000 / rdt, 'best_objective': ref_best,
                                             'note': "the reference's PrimaryAlgo(space, 'random').suggest(10) x 100, one core"}
    except ImportError as exc:
        out['config1_reference_cpu_path'] = {'unavailable': str(exc)}
    space = spaces.rosen2()
    for label, fn in (('space.sample(10)', lambda: space.sample(10)),
                      ('space.sample(1)', lambda: space.sample(1)),
                      ('points in space (10)', lambda: space.contains_points(pts)),
                      ('mixed64.sample(10)', lambda m=spaces.mixed64(): m.sample(10))):
        from orion_b200 import mt_stream
        for stream in ('reference stream', 'Philox stream'):
            if stream == 'Philox stream' and 'sample' not in label:
                continue
            mt_stream.ENABLED = stream == 'reference stream'
            fn()
            t0 = time.perf_counter()
            for _ in range(300):
                fn()
            key = 'config1_us_per_call ' + label + (' [%s]' % stream if 'sample' in label else '')
            out[key] = (time.perf_counter() - t0) / 300 * 1e6
        mt_stream.ENABLED = True
    # config 2: Space.sample 1M points, real16
    c = spaces.real16().compiled()
    for n in (1_000_000, 100_000_000):
        b = c.empty_columns(n)
        ms = timed(lambda: c.sample(n, 20240002, with_mask=True, out=b), 20 if n < 10 ** 7 else 5)
        out['config2_real16_n%d' % n] = {'ms': ms, 'candidates_per_s': n / ms * 1e3, 'GBps_written': 128 * n / ms / 1e6}
        del b
    # config 4: cat32x100 transform + reverse, 10 chunks of 1M
    c = build_required_space('real', spaces.cat32x100()).compiled()
    n = 1_000_000
    b = c.sample(n, 20240004)
    tb = c.empty_columns(n, transformed=True)
    back = c.empty_columns(n)
    t_ms = timed(lambda: c.transform(b, out=tb), 10)
    r_ms = timed(lambda: c.reverse(tb, out=back), 10)
    f_ms = timed(lambda: c.sample(n, 20240004, out=b, transformed_out=tb), 10)
    out['config4_cat32x100_per_1M_chunk'] = {
        'transform_ms': t_ms, 'reverse_ms': r_ms, 'fused_sample_transform_ms': f_ms,
        'points_per_s_transform_plus_reverse': n / (t_ms + r_ms) * 1e3,
        'transform_GBps': (128 + 25600) * n / t_ms / 1e6, 'reverse_GBps': (25600 + 128) * n / r_ms / 1e6,
        'round_trip_ok': bool(torch.equal(back.buf, b.buf))}
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
