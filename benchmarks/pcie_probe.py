"""Device -> pinned-host copy ceiling of the box, 1 .. N GPUs at once (run under gpurun;
N > 1: `python -m torch.distributed.run --nproc-per-node N benchmarks/pcie_probe.py`).

Every rank copies `--gb` GB from its GPU into its own pinned host buffer, `--reps` times,
all ranks starting together; the aggregate is what the end-to-end leg of bench.py (survivors
copied to pinned host memory every step) can reach at most on this host.  Two placements of
the pinned buffer are measured: wherever the process happens to run, and with the process
bound to the CPUs NVML names as local to its GPU before the buffer is allocated and touched
(NUMA-local).  Prints one JSON line (rank 0)."""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist


def bind_to_gpu_cpus(local_rank):
    """CPU affinity of this process := the CPUs local to its GPU (NVML); returns the count or None."""
    try:
        import pynvml
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
            handle = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
        allowed = set(os.sched_getaffinity(0))
        cpus = [c for c in cpus if c in allowed] or sorted(allowed)
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gb', type=float, default=3.2)
    ap.add_argument('--reps', type=int, default=4)
    args = ap.parse_args()
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    nbytes = int(args.gb * 1e9)
    src = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    out = {"n_gpus": world, "bytes_per_gpu_per_copy": nbytes, "reps": args.reps}
    for label in ('unbound', 'gpu_local_cpus'):
        bound = bind_to_gpu_cpus(local_rank) if label == 'gpu_local_cpus' else None
        host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        host.fill_(0)                                       # first touch from (possibly bound) this process
        host.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            host.copy_(src, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.barrier()
        wall = time.perf_counter() - t0
        out[label] = {"per_gpu_GBps_own_clock": nbytes * args.reps / (ms * 1e-3) / 1e9,
                      "aggregate_GBps": world * nbytes * args.reps / (float(t.item()) * 1e-3) / 1e9,
                      "cpus_bound": bound, "wall_s": wall}
        del host
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
