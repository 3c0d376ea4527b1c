#!/usr/bin/env python
"""Benchmark of the candidate-generation hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm (oracle port of the reference path)

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
    ap.add_argument('--e2e-steps', type=int, default=4)
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
    ap.add_argument('--exchange', default='auto', choices=['auto', 'nccl', 'peer'],
                    help='multi-GPU dedup exchange: peer-memory stores from inside the kernels, NCCL all-to-all, '
                         'or auto (peer where the ranks can map each other, else NCCL)')
    return ap.parse_args()


# --------------------------------------------------------------------------
# CPU arm: the oracle port, all host cores
# --------------------------------------------------------------------------
def _cpu_worker(job):
    """One process: the whole path on `n` candidates with numpy (oracle port)."""
    import numpy as np
    from oracle import space_oracle as so
    n, base, seed, n_hist = job
    specs = _mixed64_specs()
    hist_cols, _ = so.sample_philox(specs, n_hist, seed, base_index=HISTORY_BASE + base)
    history = set(zip(*[x.tolist() for x in so.fingerprint(hist_cols)]))
    t0 = time.perf_counter()
    # planted duplicates: 10 % history counters, 1 % repeat an earlier counter
    cols, failed = so.sample_philox(specs, n, seed, base_index=base)
    dup = np.arange(n) % 10 == 3
    for c, h in zip(cols, hist_cols):
        c[dup] = h[np.arange(n)[dup] % n_hist]
    ok = so.contains(specs, cols)
    t = so.transform(specs, 'real', cols)
    back = so.reverse(specs, 'real', t)
    lo, hi = so.fingerprint(back)
    keep = so.dedup_fingerprints(history, lo, hi)
    dt = time.perf_counter() - t0
    assert ok.all() and not failed.any() and len(keep) == n
    return dt, int(keep.sum())


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


def cpu_path_rate(n_per_proc, procs, steps=1):
    """candidates/s of the oracle port with `procs` processes; returns
    (rate, total candidates, seconds)."""
    import multiprocessing as mp
    ctx = mp.get_context('spawn')
    jobs = [(n_per_proc, p * n_per_proc, SEED, max(n_per_proc // 4, 16)) for p in range(procs)]
    total_t = 0.0
    total_n = 0
    with ctx.Pool(procs) as pool:
        pool.map(_cpu_worker, [(256, 0, SEED, 64)] * procs)          # warm imports
        for _ in range(steps):
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


def run_reference(args):
    """`--impl reference`: the reference's CPU path (numpy oracle port -- the
    reference itself is pure Python and cannot travel to the GPU box), all host
    cores, on bounded samples of the same workload."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    procs = host_cores()
    n_per_proc = args.ref_cands
    for _ in range(args.warmup and 1):
        cpu_path_rate(2000, procs)
    rate, total_n, secs = cpu_path_rate(n_per_proc, procs, steps=args.steps)
    sample = "%d candidates/step (%d per process x %d processes) of the mixed64 workload" % (
        n_per_proc * procs, n_per_proc, procs)
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": procs, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def workload_config(args, world):
    return {
        "workload": "mixed64 (16 uniform + 8 loguniform + 8 normal Real, 16 uniform Integer, 16 Categorical x 8): "
                    "sample + contains + transform('real') + reverse + fingerprint + dedup vs history",
        "candidates_per_gpu_per_step": args.cands, "history_rows_per_gpu": args.history,
        "planted_duplicates": "10% history rows, 1% earlier rows of the batch",
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


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from orion_b200 import spaces
    from orion_b200.dist import ShardedHistory
    from orion_b200.history import TrialHistory, table_slots_for
    from orion_b200.transformer import build_required_space

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback for the product path")
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node == --gpus"

    n = args.cands
    K, W = args.steps, args.warmup
    KS = min(K, 5)                                    # steps of the staged-kernels comparison pass
    KR = min(K, 20)                                   # steps of the serial per-kernel pass behind the timed region
    total_steps = K + KR + KS + W + args.e2e_steps + 2  # timed + serial pass + staged pass + e2e (+ its warm-up)
    space = spaces.mixed64()
    tspace = build_required_space('real', space)
    compiled = tspace.compiled()

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

    index_dev = [step_index(s) for s in range(W + K)]
    # two buffer sets (28 GB each) serve the pipeline, the serial pass and the e2e leg
    from orion_b200.pipeline import CandidatePipeline
    # default: ob2_suggest, ONE kernel for sample -> transform -> reverse -> fingerprint (Random.suggest fast path)
    one_kernel = not (args.staged or args.unfused or args.fp_in_sampler)
    pipe = CandidatePipeline(compiled, sharded, n, depth=2,
                             fingerprint_in='suggest' if one_kernel else 'sample' if args.fp_in_sampler else 'reverse')
    batch, tbatch, back = pipe.slots[0].batch, pipe.slots[0].tbatch, pipe.slots[0].back
    fused = not args.unfused
    # extra kernels of mine per step for the exchange: NCCL path bucket+pad+scatter; peer path send+counts+pad+return+collect
    exchange_launches = 0 if world == 1 else (4 if peer_exchange else 3)
    fp_in_reverse = fused and not args.fp_in_sampler     # fingerprints computed by the reverse kernel (default)
    rev = 'reverse+fingerprint' if fp_in_reverse else 'reverse'
    kernel_names = ['sample+transform', rev, 'dedup'] if fused else ['sample', 'transform', rev, 'dedup']
    staged_names = list(kernel_names)
    if one_kernel:
        kernel_names = ['suggest', 'dedup']
    step_launches = (3 if one_kernel else 4 if fused else 5) + exchange_launches    # dedup = 2 launches
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(kernel_names) + 1)] for _ in range(K)]
    launches = [0]

    def one_step(step, events=None, staged=False):
        j = 1
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
        compiled.reverse(tbatch, out=back, with_fingerprint=fp_in_reverse)
        if events:
            events[j + 1].record()
        sharded.dedup(back.fp if fp_in_reverse else batch.fp, n, per_rank_n=n, keep=batch.keep, check=False)
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
        # ---- timed region: K steps through the stream pipeline (sample+transform of step k+1
        # overlaps reverse and dedup of step k; every result is still materialised)
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
        kept_last = int(last_batch.keep[:n].sum().item())
        mask_ok = bool(last_batch.mask[:n].all().item())
        round_trip_ok = bool(torch.equal(last_back.buf, last_batch.buf))
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
        kept_last = int(batch.keep[:n].sum().item())
        mask_ok = bool(batch.mask[:n].all().item())
        round_trip_ok = bool(torch.equal(back.buf, batch.buf))
    if world > 1:        # did any (source, owner) lane overflow during the run?
        sharded.check_overflow() if peer_exchange else sharded.check_overflow(n)
    hist_rows1 = len(history)

    per_kernel_ms = {}
    for j, name in enumerate(kernel_names):
        per_kernel_ms[name] = sum(ev[s][j].elapsed_time(ev[s][j + 1]) for s in range(n_ev)) / n_ev

    # ---- for comparison: the same K steps through the separate kernels (what an algorithm that works in the
    # transformed space pays: the transformed columns are read back by reverse), one stream
    staged_info = None
    if one_kernel:
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

    # ---- end to end through the public API with host buffers
    e2e = None
    if args.e2e_steps > 0:
        from orion_b200.pipeline import CandidatePipeline
        E = args.e2e_steps
        host_idx = [torch.empty(n, dtype=torch.int64).pin_memory() for _ in range(E + 2)]
        for j, h in enumerate(host_idx):
            h.copy_(step_index(W + K + KR + KS + j))
        # Survivors only cross PCIe: after dedup the kept rows are gathered into a dense block on
        # the device (ob2_compact) and each column's first n_kept elements are copied to pinned
        # host memory.  Two buffer sets: compaction and copies of step j overlap step j+1.
        from orion_b200.columns import _SIZE
        ns = batch.n_stride
        # ... in the wire layout: Integer / Categorical columns at their narrowest lossless width
        # (mixed64: 288 instead of 448 bytes per candidate)
        wire = not args.e2e_natural
        out_meta = compiled.wire_meta if wire else compiled.col_meta
        out_row = compiled.wire_row_bytes if wire else compiled.row_bytes
        host_cols = [torch.empty(out_row * ns, dtype=torch.uint8).pin_memory() for _ in range(2)]
        host_cnt = [torch.zeros(1, dtype=torch.int64).pin_memory() for _ in range(2)]
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
                for units, dt in out_meta:
                    start, nbytes = units * ns, k * _SIZE[dt]
                    host_cols[b][start:start + nbytes].copy_(dense[b].buf[start:start + nbytes], non_blocking=True)
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
               "api": "CandidatePipeline.submit(counters copied from pinned host) -> "
                      + ("suggest (one kernel) -> dedup " if one_kernel else "sample+transform -> reverse || dedup ") +
                      "-> ob2_compact" + ("_wire" if wire else "") + " -> surviving candidates' columns copied to pinned host (copies overlap the next step)"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks_path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(peaks_path):
        peak = float(json.load(open(peaks_path))['hbm_gbs'])
        peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"
    algo_bytes = {'sample': BYTES_SAMPLE, 'transform': BYTES_TRANSFORM, 'reverse': BYTES_REVERSE, 'dedup': BYTES_DEDUP,
                  'sample+transform': BYTES_SAMPLE + BYTES_TRANSFORM,
                  'reverse+fingerprint': BYTES_REVERSE,
                  # the yard-stick stays the table's figure: fusion shows up as frac > naive
                  'suggest': BYTES_SAMPLE + BYTES_TRANSFORM + BYTES_REVERSE}
    algo_bytes = {k: algo_bytes[k] for k in kernel_names}
    dominant = max(per_kernel_ms, key=per_kernel_ms.get)
    achieved = algo_bytes[dominant] * n / (per_kernel_ms[dominant] * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, 'profiles', 'traffic.json')
    if os.path.exists(tpath):
        traffic = json.load(open(tpath)).get(dominant)
    kernels = {k: {"ms": per_kernel_ms[k], "algorithmic_GBps": algo_bytes[k] * n / (per_kernel_ms[k] * 1e-3) / 1e9,
                   "frac_of_hbm_peak": algo_bytes[k] * n / (per_kernel_ms[k] * 1e-3) / 1e9 / peak}
               for k in kernel_names}
    step_bytes = sum(algo_bytes.values())
    ms_per_step = elapsed_ms / K
    line = {
        "metric": METRIC, "value": world * n * K / (elapsed_ms * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, world),
        "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     # the yard-stick counts the 448 B/candidate re-read that fusing transform into the
                     # sampler removes; measured DRAM traffic (ncu, profiles/traffic.json) over the same time:
                     "traffic_frac": (traffic * n / 12_500_000 / (per_kernel_ms[dominant] * 1e-3) / 1e9 / peak
                                      if traffic else None),
                     "how": ("CUDA events around each kernel over min(K, 20) serial steps run right after the timed region "
                             "(inside it kernels of adjacent steps overlap; those contended durations are in "
                             "kernels_in_timed_region_ms)" if pipelined else
                             "CUDA events around each kernel inside the timed region"),
                     "algorithmic_bytes_per_candidate": algo_bytes[dominant]},
        "kernels": kernels,
        "step_roofline": {"algorithmic_bytes_per_candidate": step_bytes,
                          "achieved_GBps": step_bytes * n / (ms_per_step * 1e-3) / 1e9,
                          "frac": step_bytes * n / (ms_per_step * 1e-3) / 1e9 / peak},
        "e2e": e2e, "gpu_launches": timed_launches, "clocks": clock_info,
        "timed_region": ("K steps through orion_b200.pipeline.CandidatePipeline (%s, 2 buffer sets)"
                         % ("suggest and dedup on 2 streams" if one_kernel else "3 streams") if pipelined
                         else "K steps on one stream, kernels back to back"),
        "serial_ms_per_step": serial_ms,
        "kernels_in_timed_region_ms": contended_ms,
        "staged_path": staged_info,
        "history_table": {"slots": history.n_slots, "bytes": history.n_slots * 32,
                          "load_at_end": hist_rows1 / float(history.n_slots)},
        "checks": {"kept_last_step": kept_last, "kept_frac_last_step": kept_last / float(n), "all_in_space": mask_ok,
                   "reverse_of_transform_is_identity": round_trip_ok,
                   "history_rows_before": hist_rows0, "history_rows_after": hist_rows1},
    }
    line["config"]["exchange"] = ("none (1 GPU)" if world == 1 else
                                  "peer-memory inboxes over CUDA IPC (ob2_xchg_*)" if peer_exchange else "NCCL all-to-all")
    if not args.no_cpu_baseline and world == 1:
        procs = host_cores()
        n_per_proc = 200000
        rate, total_n, secs = cpu_path_rate(n_per_proc, procs, steps=1)
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": procs, "kind": "port",
                                "sample": "%d candidates (%d per process x %d processes, %.1f s) of the same workload "
                                          "through the numpy oracle" % (total_n, n_per_proc, procs, secs)}
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
