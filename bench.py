#!/usr/bin/env python
"""Benchmark of the candidate-generation hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the unmodified reference (oracle/_ref)
    python bench.py --workload cfg5 | cfg2                   # BASELINE.json configs 5 / 2 as the main line

One "step" = one pass of the whole path over one batch of synthetic candidates
of the `mixed64` space: sample (Philox, fused bounds check) ->
transform('real') -> reverse -> fingerprint -> dedup against the history shard.
By default the first four stages are ONE kernel (ob2_suggest: every stage's
result is still written to HBM, nothing is read back); `--staged` runs them as
separate kernels, and the default JSON line carries that path's timings too
(`staged_path`).  Per GPU a step processes CANDS_PER_GPU candidates against a history
shard of HISTORY_PER_GPU rows; at 8 GPUs that is exactly BASELINE.json config 5
(1e8 candidates vs 1e9-row history) with config 3's transform chain on top.
10 % of each batch are planted duplicates of history rows and 1 % duplicates of
earlier rows of the same batch (continuous dimensions never collide by chance).

Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CANDS_PER_GPU = 12_500_000
HISTORY_PER_GPU = 125_000_000
HISTORY_BASE = 1 << 40
SEED = 20240005
METRIC = "deduplicated candidates/sec (64-dim mixed space)"
UNIT = "candidates/s"

# algorithmic HBM bytes per candidate (SURVEY.md 8d / BASELINE.md section 5; fixed yard-sticks)
BYTES_SAMPLE = 448
BYTES_TRANSFORM = 448 + 1408
BYTES_REVERSE = 1408 + 448
BYTES_DEDUP = 16 + 32 + 16 + 4 + 4 + 1      # fp read, probe sector, insert, scratch w+r, keep
# what the one-kernel suggest MUST move (every API-visible result written once, nothing re-read):
# sampled 448 + transformed 1408 + handed-out 448 + fingerprint 16 + mask 1
BYTES_SUGGEST_COMPULSORY = 448 + 1408 + 448 + 16 + 1
BYTES_SAMPLE_FP_COMPULSORY = 448 + 16 + 1   # config 5's sampler: columns + fingerprint + mask
BYTES_CFG5 = 581                            # SURVEY.md 8d yard-stick for sample + fingerprint + dedup (incl. exchange staging)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--cands', type=int, default=CANDS_PER_GPU, help='candidates per GPU per step')
    ap.add_argument('--history', type=int, default=HISTORY_PER_GPU, help='history rows per GPU')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--ref-cands', type=int, default=100000,
                    help='--impl reference: candidates per process per step (a bounded sample of the workload)')
    ap.add_argument('--e2e-steps', type=int, default=8)
    ap.add_argument('--workload', default='path', choices=['path', 'cfg5', 'cfg2'],
                    help="path: the whole hot path (config 5's dedup with config 3's transform chain on top, the default "
                         "line); cfg5: sample + fingerprint + dedup only (BASELINE.json config 5 as written, 581 B/candidate "
                         "yard-stick); cfg2: Space.sample on real16.  The default line carries cfg5 and cfg2 as sub-objects.")
    ap.add_argument('--dedup-grid', type=int, default=None,
                    help="CTAs of the dedup's first pass (ob2_set_set_dedup_grid): 0 = one thread per entry; default: the "
                         "pipeline's choice (0 next to ob2_suggest, a walking grid next to config 5's sampler)")
    ap.add_argument('--no-extra', action='store_true', help='default workload only: skip the cfg5 / cfg2 / sustained sub-measurements')
    ap.add_argument('--sustain', type=float, default=2.0,
                    help='seconds of back-to-back steps for the sustained-clock sub-measurement (0 = off)')
    ap.add_argument('--ref-impl', default='auto', choices=['auto', 'literal', 'port'],
                    help='--impl reference: the unmodified reference package (oracle/_ref), the numpy port, or auto = '
                         'the reference when it is staged')
    ap.add_argument('--ref-literal-cands', type=int, default=96,
                    help='--impl reference with oracle/_ref: candidates per process per step through the unmodified reference')
    ap.add_argument('--three-buffers', action='store_true',
                    help='keep the pre-transform draw in a buffer of its own beside the handed-out rows (448 B/candidate '
                         'more to write; under the real requirement the two are equal bit for bit).  Default: ob2_suggest '
                         'with back == cols')
    ap.add_argument('--staged', action='store_true',
                    help='sample+transform, reverse+fingerprint and dedup as separate kernels (the path of an '
                         'algorithm that computes in the transformed space) instead of the one-kernel ob2_suggest')
    ap.add_argument('--unfused', action='store_true', help='sample and transform as two kernels')
    ap.add_argument('--fp-in-sampler', action='store_true',
                    help='compute the row fingerprints in the sampler instead of the reverse kernel')
    ap.add_argument('--e2e-natural', action='store_true',
                    help='e2e leg: copy the survivors out at the natural column widths (int64 / int32) instead of '
                         'the narrowed wire layout')
    ap.add_argument('--serial', action='store_true', help='no stream pipeline: one stream, kernels back to back')
    ap.add_argument('--exchange', default='auto', choices=['auto', 'nccl', 'peer', 'peer-copy'],
                    help='multi-GPU dedup exchange: peer-memory stores from inside the kernels, NCCL all-to-all, '
                         'or auto (peer where the ranks can map each other, else NCCL)')
    return ap.parse_args()


# --------------------------------------------------------------------------
# CPU arm.  Two baselines, both on all host cores (one process per core):
#   literal -- the UNMODIFIED reference (oracle/_ref, staged by oracle/make_ref.py):
#              PrimaryAlgo(space, 'random').suggest(n) = Space.sample + `assert point in space`,
#              pure Python (src/orion/core/worker/primary_algo.py:44-56); the stages the snapshot
#              lacks (transform / reverse / dedup) run through the numpy port on the points it drew;
#   port    -- the vectorised numpy restatement (oracle/space_oracle.py) of the whole path.
# --------------------------------------------------------------------------
CPU_HISTORY_ROWS = 12_500_000          # per process: sorted fingerprint array the port dedups against


def _mixed64_specs():
    from oracle.space_oracle import DimSpec
    import numpy as np
    specs = []
    for i in range(16):
        specs.append(DimSpec('u%d' % i, 'real', prior='uniform', args=(-1, 2)))
    for i in range(8):
        specs.append(DimSpec('l%d' % i, 'real', prior='reciprocal', args=(1e-5, 1)))
    for i in range(8):
        specs.append(DimSpec('g%d' % i, 'real', prior='norm', args=(0, 1)))
    for i in range(16):
        specs.append(DimSpec('i%d' % i, 'integer', prior='uniform', args=(-100, 200)))
    for i in range(16):
        specs.append(DimSpec('c%d' % i, 'categorical', categories=tuple(range(8)),
                             probs=tuple(np.tile(1. / 8, 8))))
    return specs


def _real16_specs():
    from oracle.space_oracle import DimSpec
    return ([DimSpec('u%d' % i, 'real', prior='uniform', args=(-1, 2)) for i in range(8)] +
            [DimSpec('l%d' % i, 'real', prior='reciprocal', args=(1e-5, 1)) for i in range(8)])


_CPU = {}          # per-process state of the pool workers


def _cpu_init(kind, workload, hist_rows):
    """Pool initializer (untimed): the space, and the history this process dedups against."""
    import numpy as np
    from oracle import space_oracle as so
    _CPU['workload'] = workload
    _CPU['specs'] = _real16_specs() if workload == 'cfg2' else _mixed64_specs()
    if kind == 'literal':
        from oracle import ref_harness
        space = ref_harness.build_space(ref_harness.REAL16_ARGS if workload == 'cfg2' else ref_harness.MIXED64_ARGS)
        _CPU['algo'] = ref_harness.primary_algo(space, 'random')
    if workload != 'cfg2':
        rng = np.random.RandomState(os.getpid() & 0x7FFFFFFF)
        filler = rng.randint(1, 2 ** 62, size=(hist_rows, 2), dtype=np.int64).view(np.uint64)
        _CPU['history'] = so.SortedHistory(filler[:, 0], filler[:, 1])


def _cpu_worker(job):
    """One process, one step: the whole path on `n` candidates.  Returns (seconds, kept)."""
    import numpy as np
    from oracle import space_oracle as so
    kind, n, base, seed = job
    specs, workload = _CPU['specs'], _CPU['workload']
    history = _CPU.get('history')
    if kind == 'port' and workload != 'cfg2':
        # planted duplicates need known history rows: a quarter batch of real rows joins the history (untimed)
        hist_cols, _ = so.sample_philox(specs, max(n // 4, 16), seed, base_index=HISTORY_BASE + base)
        history.insert(*so.fingerprint(hist_cols))
    t0 = time.perf_counter()
    if kind == 'literal':
        np.random.seed((seed + base) & 0x7FFFFFFF)
        points = _CPU['algo'].suggest(n)               # the reference: sample + per-point `in space`
        if workload == 'cfg2':
            return time.perf_counter() - t0, n
        cols = [np.asarray(c, dtype={'real': np.float64, 'integer': np.int64, 'categorical': np.int32}[sp.kind])
                for c, sp in zip(zip(*points), specs)]
    else:
        cols, failed = so.sample_philox(specs, n, seed, base_index=base)
        if workload == 'cfg2':
            return time.perf_counter() - t0, n
        dup = np.arange(n) % 10 == 3                   # 10 % copies of history rows
        for c, h in zip(cols, hist_cols):
            c[dup] = h[np.arange(n)[dup] % len(h)]
        ok = so.contains(specs, cols)
        assert ok.all() and not failed.any()
    if workload == 'cfg5':
        lo, hi = so.fingerprint(cols)
    else:
        t = so.transform(specs, 'real', cols)
        back = so.reverse(specs, 'real', t)
        lo, hi = so.fingerprint(back)
    keep = history.dedup(lo, hi)
    dt = time.perf_counter() - t0
    assert len(keep) == n
    return dt, int(keep.sum())


def cpu_path_rate(kind, n_per_proc, procs, steps=1, workload='path', warm=True):
    """candidates/s of the CPU path with `procs` processes; returns (rate, total candidates, seconds)."""
    import multiprocessing as mp
    ctx = mp.get_context('spawn')
    hist_rows = CPU_HISTORY_ROWS if kind == 'port' else 100_000
    total_t = 0.0
    total_n = 0
    with ctx.Pool(procs, initializer=_cpu_init, initargs=(kind, workload, hist_rows)) as pool:
        if warm:
            pool.map(_cpu_worker, [(kind, 8 if kind == 'literal' else 256, 0, SEED)] * procs)      # imports, first calls
        for step in range(steps):
            jobs = [(kind, n_per_proc, (step * procs + p) * n_per_proc, SEED) for p in range(procs)]
            t0 = time.perf_counter()
            pool.map(_cpu_worker, jobs)
            total_t += time.perf_counter() - t0
            total_n += n_per_proc * procs
    return total_n / total_t, total_n, total_t


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def reference_available():
    try:
        from oracle import ref_harness
        return ref_harness.available()
    except Exception:
        return False


def cpu_baseline_block(workload, literal_n, port_n, procs):
    """The `cpu_baseline` object: the literal reference when oracle/_ref is there (kind
    "reference"), with the vectorised port beside it; the port alone otherwise."""
    port_rate, port_total, port_s = cpu_path_rate('port', port_n, procs, workload=workload)
    port = {"value": port_rate, "unit": UNIT, "cores": procs, "kind": "port",
            "sample": "%d candidates (%d per process x %d processes, %.1f s) through the numpy oracle; each process "
                      "dedups against a sorted array of %d fingerprints" % (port_total, port_n, procs, port_s,
                                                                            CPU_HISTORY_ROWS)}
    if not reference_available():
        return port
    rate, total, secs = cpu_path_rate('literal', literal_n, procs, workload=workload)
    return {"value": rate, "unit": UNIT, "cores": procs, "kind": "reference",
            "sample": "%d candidates (%d per process x %d processes, %.1f s): the unmodified reference's "
                      "PrimaryAlgo(space, 'random').suggest(n) (oracle/_ref), then transform / reverse / dedup -- absent "
                      "from the reference -- through the numpy port on those points" % (total, literal_n, procs, secs),
            "per_core": rate / procs, "port": port}


def run_reference(args):
    """`--impl reference`: the reference's own CPU implementation of the path on all host cores
    -- the unmodified package staged in oracle/_ref when it is there, else the numpy port --
    on bounded samples of the same workload."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    procs = host_cores()
    literal = reference_available() if args.ref_impl == 'auto' else args.ref_impl == 'literal'
    kind = 'literal' if literal else 'port'
    n_per_proc = args.ref_literal_cands if literal else args.ref_cands
    rate, total_n, secs = cpu_path_rate(kind, n_per_proc, procs, steps=args.steps, workload=args.workload,
                                        warm=args.warmup > 0)
    sample = "%d candidates/step (%d per process x %d processes) of the %s workload, %s" % (
        n_per_proc * procs, n_per_proc, procs, args.workload,
        "through the unmodified reference (oracle/_ref): PrimaryAlgo.suggest; stages the snapshot lacks via the numpy port"
        if literal else "through the numpy oracle port (no reference package staged)")
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": procs, "kind": "reference" if literal else "port",
                         "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


WORKLOADS = {
    'path': "mixed64 (16 uniform + 8 loguniform + 8 normal Real, 16 uniform Integer, 16 Categorical x 8): "
            "sample + contains + transform('real') + reverse + fingerprint + dedup vs history",
    'cfg5': "mixed64: sample + contains + fingerprint + dedup vs sharded history (BASELINE.json config 5, no transform chain)",
    'cfg2': "real16 (8 uniform + 8 loguniform Real): Space.sample (BASELINE.json config 2)",
}


def workload_config(args, world):
    """Identical for the b200 and the reference arm (the driver compares the two)."""
    if args.workload == 'cfg2':
        return {"workload": WORKLOADS['cfg2'], "candidates_per_gpu_per_step": args.cands,
                "parallelism": "dp%d (candidates sharded by counter)" % world,
                "l2": "outputs larger than L2 (>= 1.6 GB written per kernel)"}
    return {
        "workload": WORKLOADS[args.workload],
        "candidates_per_gpu_per_step": args.cands, "history_rows_per_gpu": args.history,
        "planted_duplicates": "10% history rows, 1% earlier rows of the batch",
        "materialised_per_candidate": ("sampled columns, mask, fingerprint, keep" if args.workload == 'cfg5' else
                                       "transformed columns, handed-out (reversed) rows, mask, fingerprint, keep"
                                       + (", and the pre-transform draw in its own buffer" if args.three_buffers else
                                          " (the pre-transform draw shares the handed-out rows' buffer: equal bit for bit)")),
        "parallelism": "dp%d (candidates sharded by counter; history hash set sharded by fingerprint)" % world,
        "l2": "inputs larger than L2 (>= 5.6 GB touched per kernel)",
    }


# --------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------
class Clocks(object):
    """SM clock / throttle-reason sampler (NVML, every few ms) that runs in a
    thread during the timed region; falls back to an nvidia-smi subprocess."""

    def __init__(self, gpu_index, period_s=0.004):
        import threading
        self.gpu = gpu_index
        self.period = period_s
        self.rows = []
        self._stop = threading.Event()
        self._thread = None
        self._smi = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            # honour CUDA_VISIBLE_DEVICES: map the CUDA ordinal through the UUID torch reports
            import torch
            uuid = str(torch.cuda.get_device_properties(gpu_index).uuid)
            self.h = None
            for i in range(pynvml.nvmlDeviceGetCount()):
                h = pynvml.nvmlDeviceGetHandleByIndex(i)
                u = pynvml.nvmlDeviceGetUUID(h)
                u = u.decode() if isinstance(u, bytes) else u
                if uuid in u:
                    self.h = h
            if self.h is None:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self._thread = threading.Thread(target=self._loop, daemon=True)
        except Exception:
            self.nv = None
            self._smi = SmiClocks(gpu_index)

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.rows.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM),
                                  nv.nvmlDeviceGetCurrentClocksEventReasons(self.h),
                                  nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0))
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        if self._thread is not None:
            self._thread.start()
        elif self._smi is not None:
            self._smi.start()

    def stop(self):
        if self._thread is None:
            return self._smi.stop() if self._smi is not None else {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        self._stop.set()
        self._thread.join(timeout=2)
        nv = self.nv
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        sm = sorted(r[0] for r in self.rows)
        bits = 0
        for r in self.rows:
            bits |= r[1]
        names = [('hw_slowdown', nv.nvmlClocksEventReasonHwSlowdown),
                 ('hw_thermal_slowdown', nv.nvmlClocksEventReasonHwThermalSlowdown),
                 ('sw_thermal_slowdown', nv.nvmlClocksEventReasonSwThermalSlowdown),
                 ('sw_power_cap', nv.nvmlClocksEventReasonSwPowerCap),
                 ('hw_power_brake_slowdown', nv.nvmlClocksEventReasonHwPowerBrakeSlowdown)]
        return {"sm_mhz": float(sm[len(sm) // 2]), "sm_min_mhz": float(sm[0]), "sm_max_mhz": self.max_mhz,
                "reasons": [n for n, b in names if bits & b], "samples": len(sm),
                "power_w_max": max(r[2] for r in self.rows), "how": "NVML sampled every %d ms during the timed region"
                % int(self.period * 1000)}


class SmiClocks(object):
    """nvidia-smi sampler running during the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(['nvidia-smi', '-i', str(self.gpu), '--query-gpu=' + self.Q,
                                       '--format=csv,noheader,nounits', '-lms', '100'],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons, power = [], [], set(), []
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for row in self.f.read().splitlines():
            parts = [x.strip() for x in row.split(',')]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
                power.append(float(parts[3]))
            except ValueError:
                continue
            for name, val in zip(names, parts[4:8]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            sm.sort()
            out = {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                   "samples": len(sm), "power_w_max": max(power) if power else None}
        return out


def bind_to_gpu_cpus(local_rank):
    """CPU affinity of this process := the CPUs NVML names as local to its GPU, so that pinned host
    buffers allocated afterwards are first touched on that GPU's NUMA node.  Returns the number of
    CPUs, or None when NVML cannot tell."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(local_rank).uuid)
        handle = None
        for i in range(pynvml.nvmlDeviceGetCount()):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            u = pynvml.nvmlDeviceGetUUID(h)
            u = u.decode() if isinstance(u, bytes) else u
            if uuid in u:
                handle = h
        if handle is None:
            return None
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
        allowed = set(os.sched_getaffinity(0))
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return None


def d2h_ceiling(world):
    """Measured device -> pinned-host copy ceiling for `world` GPUs at once (benchmarks/pcie_probe.py,
    kept in profiles/pcie_probe.json), or None."""
    path = os.path.join(ROOT, 'profiles', 'pcie_probe.json')
    if not os.path.exists(path):
        return None
    try:
        entry = json.load(open(path)).get(str(world))
    except ValueError:
        return None
    if not isinstance(entry, dict):
        return None
    best = max(entry[k]['aggregate_GBps'] for k in ('unbound', 'gpu_local_cpus') if k in entry)
    return {"aggregate_GBps": best, "source": "profiles/pcie_probe.json (benchmarks/pcie_probe.py, %d GPUs at once)" % world}


def csrc_sha16():
    """Hash of the kernel sources: stamps every ncu-derived figure kept in profiles/."""
    import glob
    import hashlib
    h = hashlib.sha256()
    files = sorted(glob.glob(os.path.join(ROOT, 'orion_b200', 'csrc', '*.cu')) +
                   glob.glob(os.path.join(ROOT, 'orion_b200', 'csrc', '*.cuh')) +
                   [os.path.join(ROOT, 'include', 'ob2.h')])
    for path in files:
        h.update(os.path.basename(path).encode())
        h.update(open(path, 'rb').read())
    return h.hexdigest()[:16]


def measured_traffic(kernel):
    """ncu dram__bytes (read + write) per launch of `kernel` at 1.25e7 candidates, from
    profiles/traffic.json -- used only when that capture was taken from THESE kernel sources."""
    tpath = os.path.join(ROOT, 'profiles', 'traffic.json')
    if not os.path.exists(tpath):
        return None, "no profiles/traffic.json"
    data = json.load(open(tpath))
    if data.get('csrc_sha16') != csrc_sha16():
        return None, ("profiles/traffic.json was captured from other kernel sources (%s, now %s): not used"
                      % (data.get('csrc_sha16'), csrc_sha16()))
    return data.get(kernel), "ncu capture of these sources (profiles/traffic.json, %s)" % data.get('capture', '?')


def hbm_peak():
    peaks_path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(peaks_path):
        return float(json.load(open(peaks_path))['hbm_gbs']), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    return 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


def measure_cfg2(dev, reps=5, warm=2):
    """BASELINE.json config 2: Space.sample on real16, n = 1e6 (launch-overhead inclusive) and
    n = 1e8 (steady state); write-only, 128 B per candidate."""
    import torch
    from orion_b200 import spaces
    compiled = spaces.real16().compiled()
    peak, _src = hbm_peak()
    out = {}
    for n in (1_000_000, 100_000_000):
        buf = compiled.empty_columns(n)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
        for r in range(warm):
            compiled.sample(n, SEED - 3, base_index=r * n, with_mask=False, out=buf)
        torch.cuda.synchronize()
        e[0].record()
        for r in range(reps):
            compiled.sample(n, SEED - 3, base_index=(warm + r) * n, with_mask=False, out=buf)
            e[r + 1].record()
        torch.cuda.synchronize()
        ms = min(e[r].elapsed_time(e[r + 1]) for r in range(reps))
        rate = n / (ms * 1e-3)
        out['n=%d' % n] = {"ms": ms, "candidates_per_s": rate, "written_GBps": 128 * rate / 1e9,
                           "frac_of_hbm_peak": 128 * rate / 1e9 / peak}
        del buf
    torch.cuda.empty_cache()
    # the same draw over Philox4x32-7 (ob2_space_set_philox_rounds): other values, fewer instructions
    c7 = spaces.real16().compiled(philox_rounds=7)
    n = 100_000_000
    buf = c7.empty_columns(n)
    e = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    c7.sample(n, SEED - 3, with_mask=False, out=buf)
    torch.cuda.synchronize()
    e[0].record()
    for r in range(reps):
        c7.sample(n, SEED - 3, base_index=(1 + r) * n, with_mask=False, out=buf)
        e[r + 1].record()
    torch.cuda.synchronize()
    ms = min(e[r].elapsed_time(e[r + 1]) for r in range(reps))
    out['n=100000000, 7 Philox rounds'] = {"ms": ms, "candidates_per_s": n / (ms * 1e-3),
                                           "frac_of_hbm_peak": 128 * n / (ms * 1e-3) / 1e9 / peak}
    del buf
    torch.cuda.empty_cache()
    out["algorithmic_bytes_per_candidate"] = 128
    out["roofline_candidates_per_s"] = peak * 1e9 / 128
    return out


def run_cfg2(args):
    """`--workload cfg2` as the main line: one step = Space.sample of `--cands` real16 candidates."""
    import torch
    import torch.distributed as dist
    from orion_b200 import spaces
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    n, K, W = args.cands, args.steps, args.warmup
    compiled = spaces.real16().compiled()
    bufs = [compiled.empty_columns(n) for _ in range(2)]
    for s in range(W):
        compiled.sample(n, SEED - 3, base_index=(s * world + rank) * n, with_mask=False, out=bufs[s % 2])
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = Clocks(local_rank)
    if rank == 0:
        clocks.start()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for s in range(K):
        compiled.sample(n, SEED - 3, base_index=((W + s) * world + rank) * n, with_mask=False, out=bufs[s % 2])
    t1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clock_info = clocks.stop() if rank == 0 else None
    ms = t0.elapsed_time(t1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # end to end: Space.sample_columns -> pinned host (the public API with a host result)
    E = max(args.e2e_steps, 1)
    host = torch.empty(bufs[0].buf.numel(), dtype=torch.uint8).pin_memory()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for j in range(E):
        compiled.sample(n, SEED - 4, base_index=(j * world + rank) * n, with_mask=False, out=bufs[0])
        host.copy_(bufs[0].buf, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = hbm_peak()
    rate1 = n * K / (ms * 1e-3)
    line = {"metric": "Space.sample candidates/sec (real16)", "value": world * rate1, "unit": UNIT, "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, world),
            "roofline": {"bound": "hbm", "kernel": "sample", "achieved": 128 * rate1 / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": 128 * rate1 / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_candidate": 128},
            "e2e": {"value": world * n * E / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": int(bufs[0].buf.numel()), "steps": E},
            "gpu_launches": K, "clocks": clock_info}
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline_block('cfg2', 256, 200000, host_cores())
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_b200(args):
    import torch
    import torch.distributed as dist
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory, table_slots_for
    from orion_b200.transformer import build_required_space

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback for the product path")
    if args.workload == 'cfg2':
        return run_cfg2(args)
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node == --gpus"

    cfg5_main = args.workload == 'cfg5'
    extras = not args.no_extra and not cfg5_main
    n = args.cands
    K, W = args.steps, args.warmup
    KS = 0 if cfg5_main else min(K, 5)                # steps of the staged-kernels comparison pass
    KR = min(K, 20)                                   # steps of the serial per-kernel pass behind the timed region
    K5 = min(K, 10) if extras else 0                  # steps of the config-5 sub-measurement
    total_steps = K + KR + KS + 3 * K5 + W + args.e2e_steps + 7  # timed + serial + staged + cfg5 (serial, pipelined) + e2e (+ warm-ups)
    space = spaces.mixed64()
    tspace = build_required_space('real', space)
    compiled = space.compiled() if cfg5_main else tspace.compiled()
    plain = space.compiled()                          # no transform chain: config 5's sampler

    # ---- history shard: rows from counters [2^40 + rank*H, ...), fingerprints only
    H = args.history
    cap_rows = H + int(0.92 * n) * total_steps + n
    if table_slots_for(cap_rows, load=0.75) > (1 << 31):
        raise SystemExit("bench.py: %d steps of %d candidates on top of %d history rows need more than the 2^31 slots "
                         "one history table holds; use fewer --steps or --cands" % (total_steps, n, H))
    # linear probing: short chains (one 128-byte line) while the table stays under ~40 % full; take the
    # roomier size when it fits the 2^31-slot limit (69 GB of the 180 GB), else settle for <= 75 %
    load = 0.4 if table_slots_for(cap_rows, load=0.4) <= (1 << 31) else 0.75
    history = TrialHistory(compiled, n_slots=table_slots_for(cap_rows, load=load))
    # 'auto': peer-memory inboxes (kernels store straight into the owners' HBM over NVLink) where the
    # ranks can map each other's memory, NCCL all-to-all otherwise
    from orion_b200.dist import PeerShardedHistory, make_sharded_history
    sharded = make_sharded_history(history, max_n=n, exchange=args.exchange)
    peer_exchange = isinstance(sharded, PeerShardedHistory)
    chunk = n
    hb = compiled.empty_columns(chunk)
    done = 0
    while done < H:
        m = min(chunk, H - done)
        b = compiled.sample(m, SEED, base_index=HISTORY_BASE + rank * H + done, with_mask=False,
                            with_fingerprint=True, out=hb if m == chunk else None)
        sharded.insert(b.fp, m)
        done += m
    torch.cuda.synchronize()
    del hb, b
    hist_rows0 = len(history)

    # ---- per-step candidate counters with planted duplicates (synthetic input, built before timing)
    def step_index(step):
        g = torch.Generator(device=dev)
        g.manual_seed(SEED + 1000 * step + rank)
        base = (step * world + rank) * n
        idx = torch.arange(base, base + n, dtype=torch.int64, device=dev)
        r = torch.rand(n, generator=g, device=dev)
        hist_pick = torch.randint(0, H * world, (n,), generator=g, device=dev, dtype=torch.int64) + HISTORY_BASE
        idx = torch.where(r < 0.10, hist_pick, idx)
        early = base + torch.randint(0, max(n // 100, 1), (n,), generator=g, device=dev, dtype=torch.int64)
        idx = torch.where((r >= 0.10) & (r < 0.11), early, idx)
        return idx

    def expected_keep(idx):
        """The keep mask this step MUST produce, from the Philox counters alone (a candidate is a
        function of its counter; continuous dimensions never collide otherwise): not a history
        counter, and the first global position (rank order) holding that counter.  Fresh counters
        of different steps are disjoint by construction."""
        if world > 1:
            everyone = torch.empty(world * n, dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(everyone, idx.contiguous())
        else:
            everyone = idx
        order = torch.argsort(everyone, stable=True)
        srt = everyone[order]
        first_sorted = torch.ones(everyone.numel(), dtype=torch.bool, device=dev)
        first_sorted[1:] = srt[1:] != srt[:-1]
        first = torch.empty_like(first_sorted)
        first[order] = first_sorted
        want = first & (everyone < HISTORY_BASE)
        return want[rank * n:(rank + 1) * n]

    def all_ranks_true(flag):
        if world > 1:
            t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            return bool(t.item())
        return bool(flag)

    index_dev = [step_index(s) for s in range(W + K)]
    # two buffer sets (28 GB each) serve the pipeline, the serial pass and the e2e leg
    from orion_b200.pipeline import CandidatePipeline
    # default: ob2_suggest, ONE kernel for sample -> transform -> reverse -> fingerprint (Random.suggest fast path)
    one_kernel = not (args.staged or args.unfused or args.fp_in_sampler) and not cfg5_main
    # Random.suggest behind PrimaryAlgo hands out the reversed rows and lets the algorithm see the transformed
    # ones; the draw before the transform is an intermediate.  By default it therefore shares its buffer with
    # the handed-out rows (ob2_suggest, back == cols); --three-buffers materialises it separately as well.
    alias = one_kernel and not args.three_buffers
    pipe = CandidatePipeline(compiled, sharded, n, depth=2, alias_back=alias, dedup_grid=args.dedup_grid,
                             fingerprint_in='sample_only' if cfg5_main else 'suggest' if one_kernel
                             else 'sample' if args.fp_in_sampler else 'reverse')
    batch, tbatch, back = pipe.slots[0].batch, pipe.slots[0].tbatch, pipe.slots[0].back
    fused = not args.unfused
    # extra kernels of mine per step for the exchange: NCCL path bucket+pad+scatter; peer path send+counts+flag+pad+collect
    exchange_launches = 0 if world == 1 else (4 if peer_exchange else 3)    # peer: send + counts + flag + collect (or bucket + pad + counts + flag, scatter)
    fp_in_reverse = fused and not args.fp_in_sampler     # fingerprints computed by the reverse kernel (default)
    rev = 'reverse+fingerprint' if fp_in_reverse else 'reverse'
    kernel_names = ['sample+transform', rev, 'dedup'] if fused else ['sample', 'transform', rev, 'dedup']
    staged_names = list(kernel_names)
    if one_kernel:
        kernel_names = ['suggest', 'dedup']
    if cfg5_main:
        kernel_names = ['sample+fingerprint', 'dedup']
    step_launches = (3 if (one_kernel or cfg5_main) else 4 if fused else 5) + exchange_launches    # dedup = 2 launches
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(kernel_names) + 1)] for _ in range(K)]
    launches = [0]
    stage = {'back': back}          # where the staged comparison pass reverses into (a buffer of its own if `alias`)

    def cfg5_step(idx, events=None):
        """BASELINE.json config 5 as written: sample (fused contains + fingerprint) -> dedup."""
        plain.sample(n, SEED, index=idx, with_mask=True, with_fingerprint=True, out=batch)
        if events:
            events[1].record()
        sharded.dedup(batch.fp, n, per_rank_n=n, keep=batch.keep, check=False)
        if events:
            events[2].record()

    def one_step(step, events=None, staged=False):
        j = 1
        if cfg5_main:
            cfg5_step(index_dev[step], events)
            launches[0] += step_launches
            return
        if one_kernel and not staged:
            compiled.suggest(n, SEED, index=index_dev[step], out=batch, transformed_out=tbatch, back_out=back)
            if events:
                events[1].record()
            sharded.dedup(back.fp, n, per_rank_n=n, keep=batch.keep, check=False)
            if events:
                events[2].record()
            launches[0] += step_launches
            return
        if fused:       # TransformedSpace.sample: one kernel draws and transforms
            compiled.sample(n, SEED, index=index_dev[step], with_mask=True, with_fingerprint=not fp_in_reverse,
                            out=batch, transformed_out=tbatch)
        else:
            compiled.sample(n, SEED, index=index_dev[step], with_mask=True, with_fingerprint=True, out=batch)
            if events:
                events[j].record()
            j += 1
            compiled.transform(batch, out=tbatch)
        if events:
            events[j].record()
        sback = stage['back']
        compiled.reverse(tbatch, out=sback, with_fingerprint=fp_in_reverse)
        if events:
            events[j + 1].record()
        sharded.dedup(sback.fp if fp_in_reverse else batch.fp, n, per_rank_n=n, keep=batch.keep, check=False)
        if events:
            events[j + 2].record()
        launches[0] += (4 if fused else 5) + exchange_launches

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    pipelined = fused and not args.serial
    n_ev = K                      # steps whose per-kernel events are averaged
    t_begin = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    clocks = Clocks(local_rank)
    contended_ms = None
    if pipelined:
        # ---- timed region: K steps through the stream pipeline (the sampling kernel of step k+1
        # overlaps dedup of step k; every result is still materialised)
        for s in range(W):
            pipe.submit(SEED, index=index_dev[s])
        pipe.drain()
        sync_all()
        pipe.timed = True
        if rank == 0:
            clocks.start()
        t_begin.record()
        for s in range(K):
            slot = pipe.submit(SEED, index=index_dev[W + s])
        pipe.drain()
        t_end.record()
        sync_all()
        clock_info = clocks.stop() if rank == 0 else None
        elapsed_ms = max_over_ranks(t_begin.elapsed_time(t_end))
        timed_launches = K * step_launches
        contended_ms = pipe.kernel_ms()
        pipe.timed = False
        last_batch, last_back = slot.batch, slot.back
    else:
        for s in range(W):
            one_step(s)
        sync_all()
        launches[0] = 0
        if rank == 0:
            clocks.start()
        t_begin.record()
        for s in range(K):
            ev[s][0].record()
            one_step(W + s, ev[s])
        t_end.record()
        sync_all()
        clock_info = clocks.stop() if rank == 0 else None
        elapsed_ms = max_over_ranks(t_begin.elapsed_time(t_end))
        timed_launches = launches[0]
        serial_ms = elapsed_ms / K
        last_batch, last_back = batch, back
    # ---- checks on the last timed step (outside the timed region)
    kept_last = int(last_batch.keep[:n].sum().item())
    mask_ok = bool(last_batch.mask[:n].all().item())
    round_trip_ok = None if (cfg5_main or alias) else bool(torch.equal(last_back.buf, last_batch.buf))
    want = expected_keep(index_dev[W + K - 1])
    keep_mask_exact = all_ranks_true(bool(torch.equal(last_batch.keep[:n].bool(), want)))
    del want
    dedup_grid_timed = pipe.dedup_grid if pipelined else 0
    sharded.set_dedup_grid(0)       # the serial passes time every kernel alone: the dedup with its one-thread-per-entry grid
    if pipelined:
        # ---- the same steps (at most 20), one stream, no overlap: clean per-kernel durations for the roofline
        n_ev = KR
        for s in range(KR):
            index_dev[W + s] = step_index(W + K + s)      # fresh counters, same duplicate mix
        sync_all()
        for s in range(KR):
            ev[s][0].record()
            one_step(W + s, ev[s])
        sync_all()
        serial_ms = sum(ev[s][0].elapsed_time(ev[s][-1]) for s in range(KR)) / KR
    if world > 1:        # did any (source, owner) lane overflow / any shard fill up during the run?
        sharded.check_overflow() if peer_exchange else sharded.check_overflow(n)
    hist_rows1 = len(history)

    per_kernel_ms = {}
    for j, name in enumerate(kernel_names):
        per_kernel_ms[name] = sum(ev[s][j].elapsed_time(ev[s][j + 1]) for s in range(n_ev)) / n_ev

    # ---- for comparison: the same K steps through the separate kernels (what an algorithm that works in the
    # transformed space pays: the transformed columns are read back by reverse), one stream
    staged_info = None
    three_info = None
    if one_kernel:
        if alias:       # the staged kernels (and the three-buffer suggest) need the separate buffer after all
            stage['back'] = compiled.empty_columns(n)
            last_rows = batch.buf.clone()
            last_serial = (W + KR - 1) if pipelined else (W + K - 1)
            compiled.suggest(n, SEED, index=index_dev[last_serial], out=batch, transformed_out=tbatch, back_out=stage['back'])
            # the one-buffer suggest wrote the same handed-out rows as the three-buffer one, and draw == handed-out
            round_trip_ok = bool(torch.equal(stage['back'].buf, last_rows)) and bool(torch.equal(stage['back'].buf, batch.buf))
            del last_rows
            tev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            for j in range(3):
                tev[j].record()
                compiled.suggest(n, SEED, index=index_dev[W], out=batch, transformed_out=tbatch, back_out=stage['back'])
            tev[3].record()
            torch.cuda.synchronize()
            three_info = {"ms": min(tev[j].elapsed_time(tev[j + 1]) for j in range(3)),
                          "bytes_per_candidate": BYTES_SUGGEST_COMPULSORY,
                          "note": "ob2_suggest with the pre-transform draw in a buffer of its own (--three-buffers)"}
        sev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(staged_names) + 1)] for _ in range(KS)]
        for s in range(KS):
            index_dev[W + s] = step_index(W + K + KR + s)
        sync_all()
        for s in range(KS):
            sev[s][0].record()
            one_step(W + s, sev[s], staged=True)
        sync_all()
        staged_info = {"steps": KS, "ms_per_step": sum(e[0].elapsed_time(e[-1]) for e in sev) / KS,
                       "kernels_ms": {name: sum(e[j].elapsed_time(e[j + 1]) for e in sev) / KS
                                      for j, name in enumerate(staged_names)},
                       "note": "same workload through ob2_sample_transform -> ob2_reverse_fingerprint -> ob2_set_dedup, "
                               "one stream, run after the timed region"}

    peak, peak_src = hbm_peak()

    # ---- BASELINE.json config 5 on its own yard-stick: sample + fingerprint + dedup, no transform chain
    cfg5_info = None
    if K5:
        base5 = W + K + KR + KS
        c5 = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K5)]
        cfg5_step(step_index(base5))                                  # warm-up (its rows become history)
        idx5 = None
        for s in range(K5):
            idx5 = step_index(base5 + 1 + s)
            c5[s][0].record()
            cfg5_step(idx5, c5[s])
        sync_all()
        c5_ms = max_over_ranks(sum(e[0].elapsed_time(e[2]) for e in c5) / K5)
        c5_want = expected_keep(idx5)
        c5_exact = all_ranks_true(bool(torch.equal(batch.keep[:n].bool(), c5_want)))
        del c5_want, idx5
        sample_ms = sum(e[0].elapsed_time(e[1]) for e in c5) / K5
        # ... and as it is meant to run: batch after batch through the pipeline, the dedup of batch k on a second
        # stream as a SMALL walking grid (ob2_set_set_dedup_grid) that shares the SMs with the issue-bound
        # sampler of batch k + 1 (with its one-thread-per-entry grid the two kernels take turns and the times add up)
        pipe5 = CandidatePipeline(plain, sharded, n, depth=2, fingerprint_in='sample_only')
        base5p = base5 + 1 + K5
        KP5 = 2 * K5                                                  # (enough batches to amortise the pipeline's fill and tail)
        idx5p = [step_index(base5p + s) for s in range(KP5 + 2)]
        for s in range(2):
            pipe5.submit(SEED, index=idx5p[s])
        pipe5.drain()
        sync_all()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        slot5 = None
        for s in range(KP5):
            slot5 = pipe5.submit(SEED, index=idx5p[2 + s])
        pipe5.drain()
        p1.record()
        sync_all()
        c5_pipe_ms = max_over_ranks(p0.elapsed_time(p1) / KP5)
        c5_want = expected_keep(idx5p[-1])
        c5_exact = c5_exact and all_ranks_true(bool(torch.equal(slot5.batch.keep[:n].bool(), c5_want)))
        sharded.set_dedup_grid(0)
        dedup_grid5 = pipe5.dedup_grid
        del c5_want, idx5p, pipe5, slot5
        torch.cuda.empty_cache()
        rate5 = n / (c5_pipe_ms * 1e-3)
        cfg5_info = {"workload": WORKLOADS['cfg5'], "steps": KP5, "ms_per_step": c5_pipe_ms, "value": world * rate5, "unit": UNIT,
                     "serial_ms_per_step": c5_ms, "dedup_grid_ctas": dedup_grid5,
                     "kernels_ms": {"sample+fingerprint": sample_ms,
                                    "dedup": sum(e[1].elapsed_time(e[2]) for e in c5) / K5},
                     "roofline": {"yardstick_bytes_per_candidate": BYTES_CFG5,
                                  "roofline_candidates_per_s_per_gpu": peak * 1e9 / BYTES_CFG5,
                                  "frac": BYTES_CFG5 * rate5 / 1e9 / peak,
                                  "frac_serial": BYTES_CFG5 * n / (c5_ms * 1e-3) / 1e9 / peak,
                                  "sampler_frac_of_hbm_peak": BYTES_SAMPLE_FP_COMPULSORY * n / (sample_ms * 1e-3) / 1e9 / peak,
                                  "note": "per GPU; ms_per_step: 2 K5 batches through CandidatePipeline(fingerprint_in="
                                          "'sample_only') on two streams, the dedup as a walking grid beside the next "
                                          "batch's sampler; kernels_ms / serial_ms_per_step: one stream, CUDA events around "
                                          "each kernel, one-thread-per-entry dedup"},
                     "keep_mask_exact": c5_exact}
    K5_used = 3 * K5 + 3 if K5 else 0        # serial: 1 + K5 steps; pipelined: 2 + 2 K5

    cfg2_info = None
    if extras and world == 1:
        cfg2_info = measure_cfg2(dev)

    # ---- sustained clocks: back-to-back steps for >= S seconds.  The table cannot take S seconds of
    # fresh rows (2^31 slots), so these steps replay counters already seen: the sampling kernel does its
    # full work, dedup finds every row (probe only).
    sustained = None
    if extras and args.sustain > 0:
        sclocks = Clocks(local_rank, period_s=0.02)
        per = max(serial_ms, 1.0)
        n_sus = int(args.sustain * 1e3 / per) + 1
        sync_all()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if rank == 0:
            sclocks.start()
        s0.record()
        for s in range(n_sus):
            one_step(W + (s % K))
        s1.record()
        sync_all()
        sus_clock = sclocks.stop() if rank == 0 else None
        sus_ms = max_over_ranks(s0.elapsed_time(s1))
        sustained = {"seconds": sus_ms * 1e-3, "steps": n_sus, "ms_per_step": sus_ms / n_sus,
                     "value": world * n * n_sus / (sus_ms * 1e-3), "unit": UNIT, "clocks": sus_clock,
                     "note": "replayed counters: the sampling kernel in full, dedup probe-only (every row already in the history)"}

    # ---- end to end through the public API with host buffers
    e2e = None
    if args.e2e_steps > 0:
        E = args.e2e_steps
        affinity0 = os.sched_getaffinity(0)
        cpus_bound = bind_to_gpu_cpus(local_rank)          # pinned buffers below: first touched NUMA-local to this GPU
        host_idx = [torch.empty(n, dtype=torch.int64).pin_memory() for _ in range(E + 2)]
        for j, h in enumerate(host_idx):
            h.copy_(step_index(W + K + KR + KS + K5_used + j))
        # Survivors only cross PCIe: after dedup the kept rows are gathered into a dense block on
        # the device (ob2_compact) and each column's first n_kept elements are copied to pinned
        # host memory.  Two buffer sets: compaction and copies of step j overlap step j+1.
        ns = batch.n_stride
        # ... in the wire layout: Integer / Categorical columns at their narrowest lossless width
        # (mixed64: 288 instead of 448 bytes per candidate)
        wire = not args.e2e_natural
        out_meta = compiled.wire_meta if wire else compiled.col_meta
        out_row = compiled.wire_row_bytes if wire else compiled.row_bytes
        host_cols = [torch.empty(out_row * ns, dtype=torch.uint8).pin_memory() for _ in range(2)]
        host_cnt = [torch.zeros(1, dtype=torch.int64).pin_memory() for _ in range(2)]
        for h in host_cols:
            h.zero_()                                      # first touch while bound
        os.sched_setaffinity(0, affinity0)                 # (the CPU baseline later wants every core again)
        dense = [compiled.empty_columns(n, wire=wire) for _ in range(2)]
        cnt_dev = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(2)]
        cscratch = [torch.empty(max(int(compiled.lib.ob2_compact_scratch_bytes(n)), 8), dtype=torch.uint8, device=dev)
                    for _ in range(2)]
        idx_dev = [torch.empty(n, dtype=torch.int64, device=dev) for _ in range(2)]
        compacted = [torch.cuda.Event() for _ in range(2)]
        copied = [torch.cuda.Event() for _ in range(2)]
        s_comp = torch.cuda.Stream(device=dev)
        s_copy = torch.cuda.Stream(device=dev)
        epipe = pipe
        epipe.drain()
        cur = torch.cuda.current_stream()
        state = {'pending': None, 'd2h': 0, 'kept': 0, 'used': [False, False]}

        def copy_out(b):
            compacted[b].synchronize()              # the host needs n_kept; the GPU already works on the next step
            k = int(host_cnt[b][0])
            with torch.cuda.stream(s_copy):
                # the survivors' rows of every column: ob2_columns_to_host, one 2-D copy per run of equal-width columns
                dense[b].copy_rows_to_host(host_cols[b], k)
                copied[b].record()
            state['d2h'] += k * out_row + 8
            state['kept'] = k

        def e2e_step(j):
            b = j % 2
            slot = epipe.slots[epipe.k % 2]
            if slot.busy:
                cur.wait_event(slot.sampled)                                    # idx_dev[b] no longer read
            idx_dev[b].copy_(host_idx[j], non_blocking=True)                    # H2D: this step's input
            slot = epipe.submit(SEED, index=idx_dev[b])
            with torch.cuda.stream(s_comp):
                s_comp.wait_event(slot.reversed)
                s_comp.wait_event(slot.deduped)
                if state['used'][b]:
                    s_comp.wait_event(copied[b])                                # dense[b] has been copied out
                slot.back.compact(keep=slot.batch.keep, out=dense[b], count=cnt_dev[b], scratch=cscratch[b], wire=wire)
                host_cnt[b].copy_(cnt_dev[b], non_blocking=True)                # D2H: how many survived
                compacted[b].record()
            state['used'][b] = True
            slot.hold = compacted[b]                                            # slot buffers are free once gathered
            if state['pending'] is not None:
                copy_out(state['pending'])                                      # D2H: the previous step's survivors
            state['pending'] = b

        def e2e_finish():
            if state['pending'] is not None:
                copy_out(state['pending'])
                state['pending'] = None
            epipe.drain()
            cur.wait_stream(s_comp)
            cur.wait_stream(s_copy)

        e2e_step(0)
        e2e_step(1)
        e2e_finish()
        sync_all()
        state['d2h'] = 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for j in range(E):
            e2e_step(2 + j)
        e2e_finish()
        e1.record()
        sync_all()
        e2e_ms = max_over_ranks(e0.elapsed_time(e1))
        e2e = {"value": world * n * args.e2e_steps / (e2e_ms * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(n * 8),
               "d2h_bytes_per_step": int(state['d2h'] // max(E, 1)),
               "steps": args.e2e_steps, "kept_last_step": state['kept'],
               "d2h_GBps_per_gpu": state['d2h'] / (e2e_ms * 1e-3) / 1e9,
               "d2h_ceiling": d2h_ceiling(world), "host_cpus_bound_per_rank": cpus_bound,
               "api": "CandidatePipeline.submit(counters copied from pinned host) -> "
                      + ("sample+fingerprint -> dedup " if cfg5_main else "suggest (one kernel) -> dedup " if one_kernel
                         else "sample+transform -> reverse || dedup ") +
                      "-> ob2_compact" + ("_wire" if wire else "") + " -> surviving candidates' columns copied to pinned host (copies overlap the next step)"}

    if e2e and e2e.get("d2h_ceiling"):
        e2e["frac_of_d2h_ceiling"] = world * e2e["d2h_GBps_per_gpu"] / e2e["d2h_ceiling"]["aggregate_GBps"]
        if e2e["frac_of_d2h_ceiling"] > 1.05:       # the ceiling was probed on ANOTHER box of the pool
            e2e["d2h_ceiling"]["note"] = ("this box moves more than the probed one (PCIe topology differs between "
                                          "the pool's boxes): re-run benchmarks/pcie_probe.py here for its own ceiling")
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # yard-stick bytes (every stage's output AND the re-reads a staged path pays) and the
    # bytes a kernel must move as built (each API-visible result written once, inputs read once)
    yard = {'sample': BYTES_SAMPLE, 'transform': BYTES_TRANSFORM, 'reverse': BYTES_REVERSE, 'dedup': BYTES_DEDUP,
            'sample+transform': BYTES_SAMPLE + BYTES_TRANSFORM, 'reverse+fingerprint': BYTES_REVERSE,
            'suggest': BYTES_SAMPLE + BYTES_TRANSFORM + BYTES_REVERSE, 'sample+fingerprint': BYTES_SAMPLE}
    compulsory = dict(yard, **{'suggest': BYTES_SUGGEST_COMPULSORY - (448 if alias else 0),
                               'sample+fingerprint': BYTES_SAMPLE_FP_COMPULSORY,
                               'sample+transform': 448 + 1408 + 1, 'reverse+fingerprint': 1408 + 448 + 16})
    yard = {k: yard[k] for k in kernel_names}
    dominant = max(per_kernel_ms, key=per_kernel_ms.get)
    dom_s = per_kernel_ms[dominant] * 1e-3
    achieved = compulsory[dominant] * n / dom_s / 1e9
    traffic, traffic_src = measured_traffic(dominant)
    if traffic is not None:
        traffic = traffic * n / 12_500_000
    kernels = {k: {"ms": per_kernel_ms[k], "algorithmic_GBps": compulsory[k] * n / (per_kernel_ms[k] * 1e-3) / 1e9,
                   "frac_of_hbm_peak": compulsory[k] * n / (per_kernel_ms[k] * 1e-3) / 1e9 / peak,
                   "frac_yardstick": yard[k] * n / (per_kernel_ms[k] * 1e-3) / 1e9 / peak}
               for k in kernel_names}
    step_bytes = BYTES_CFG5 if cfg5_main else sum(yard.values())
    ms_per_step = elapsed_ms / K
    line = {
        "metric": METRIC, "value": world * n * K / (elapsed_ms * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, world),
        "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_candidate": compulsory[dominant],
                     "algorithmic_bytes_note": "what this kernel must move: every API-visible result written once, nothing "
                                               "re-read (suggest: transformed 1408 + handed-out 448 + fingerprint 16 + mask 1"
                                               + (")" if alias else " + the pre-transform draw 448)"),
                     # SURVEY.md 8d's per-stage figure, which also counts the re-reads a staged path pays
                     # (fusion removed them, hence > 1 for the fused kernel): kept beside, never as `frac`
                     "yardstick_bytes_per_candidate": yard[dominant],
                     "frac_yardstick": yard[dominant] * n / dom_s / 1e9 / peak,
                     "traffic_frac": traffic / dom_s / 1e9 / peak if traffic else None,
                     "traffic_source": traffic_src,
                     "how": ("CUDA events around each kernel over min(K, 20) serial steps run right after the timed region "
                             "(inside it kernels of adjacent steps overlap; those contended durations are in "
                             "kernels_in_timed_region_ms)" if pipelined else
                             "CUDA events around each kernel inside the timed region")},
        "kernels": kernels,
        "step_roofline": {"yardstick_bytes_per_candidate": step_bytes,
                          "frac_yardstick": step_bytes * n / (ms_per_step * 1e-3) / 1e9 / peak,
                          "note": "per GPU, whole step, SURVEY.md 8d yard-stick (a fused path scores above a staged one)"},
        "e2e": e2e, "gpu_launches": timed_launches, "clocks": clock_info,
        "timed_region": ("K steps through orion_b200.pipeline.CandidatePipeline (%s, 2 buffer sets)"
                         % ("one sampling kernel and dedup on 2 streams" if (one_kernel or cfg5_main) else "3 streams")
                         if pipelined else "K steps on one stream, kernels back to back"),
        "serial_ms_per_step": serial_ms,
        "kernels_in_timed_region_ms": contended_ms,
        "staged_path": staged_info, "suggest_three_buffers": three_info,
        "dedup_grid_ctas": dedup_grid_timed,
        "cfg5": cfg5_info, "cfg2": cfg2_info, "sustained": sustained,
        "history_table": {"slots": history.n_slots, "bytes": history.n_slots * 32,
                          "load_at_end": hist_rows1 / float(history.n_slots)},
        "checks": {"keep_mask_exact": keep_mask_exact,
                   "keep_mask_exact_how": "last timed step, every rank: keep == (first global position of its Philox "
                                          "counter) & (not a history counter), from the all-gathered counters",
                   "kept_last_step": kept_last, "kept_frac_last_step": kept_last / float(n), "all_in_space": mask_ok,
                   "reverse_of_transform_is_identity": round_trip_ok,
                   "history_rows_before": hist_rows0, "history_rows_after": hist_rows1},
        "exchange": ("none (1 GPU)" if world == 1 else
                     ("peer-memory inboxes over CUDA IPC (ob2_xchg_*), %s" % ("DMA transfers" if sharded.copy else "remote stores from the kernels"))
                     if peer_exchange else "NCCL all-to-all"),
        "kernel_sources_sha16": csrc_sha16(),
    }
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline_block(args.workload, 256, 200000, host_cores())
    emit(line)
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line, on the process's real stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    args = parse_args()
    # Libraries chat on stdout (NCCL prints its version banner there): park fd 1 on stderr for
    # the duration of the run so that stdout carries exactly the JSON line.
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
