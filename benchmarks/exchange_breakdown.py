#!/usr/bin/env python
"""Where does the multi-GPU dedup time go?  CUDA events around each phase of
PeerShardedHistory.dedup (send | barrier | pad+dedup | barrier | collect), max over ranks.
  torchrun --nproc-per-node N benchmarks/exchange_breakdown.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from orion_b200 import spaces
from orion_b200.dist import make_sharded_history
from orion_b200.history import TrialHistory, table_slots_for
from orion_b200.transformer import build_required_space

rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', 0)))
dev = torch.device('cuda', torch.cuda.current_device())
if world > 1:
    dist.init_process_group('nccl', device_id=dev)
n, steps = 12_500_000, 8
compiled = build_required_space('real', spaces.mixed64()).compiled()
history = TrialHistory(compiled, n_slots=table_slots_for(n * (steps + 4), load=0.5))
sh = make_sharded_history(history, max_n=n, exchange='peer' if world > 1 else 'auto')
batch = compiled.empty_columns(n)
keep = torch.empty(batch.n_stride, dtype=torch.uint8, device=dev)
names = ['send', 'barrier1', 'pad+dedup', 'barrier2', 'collect']
acc = dict.fromkeys(names, 0.0)
for s in range(steps + 2):
    compiled.sample(n, 5, base_index=(s * world + rank) * n, with_mask=False, with_fingerprint=True, out=batch)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    ev[0].record()
    if world > 1:
        sh.xchg.send(batch.fp, n, rank * n)
        ev[1].record()
        sh._barrier()
        ev[2].record()
        sh.xchg.dedup(history)
        ev[3].record()
        sh._barrier()
        ev[4].record()
        sh.xchg.collect(n, keep)
    else:
        for e in ev[1:3]:
            e.record()
        history.dedup_fingerprints(batch.fp, n, keep=keep)
        for e in ev[3:5]:
            e.record()
    ev[5].record()
    torch.cuda.synchronize()
    if s >= 2:
        for j, name in enumerate(names):
            acc[name] += ev[j].elapsed_time(ev[j + 1]) / steps
t = torch.tensor([acc[k] for k in names], dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print('world %d: ' % world + '  '.join('%s %.3f' % (k, v) for k, v in zip(names, t.tolist())) +
          '  total %.3f ms' % float(t.sum()))
if world > 1:
    dist.destroy_process_group()
