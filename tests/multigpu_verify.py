#!/usr/bin/env python
"""Exactness of the sharded dedup on REAL GPUs (one process per GPU, torchrun).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tests/multigpu_verify.py [--cands 12500000] [--history 25000000] \
        [--steps 3] [--exchange both] [--out gpurun_out/verify_N.json]

For each exchange (``peer`` = CUDA-IPC inboxes written by the kernels over
NVLink, ``nccl`` = all-to-all) the ranks run ``--steps`` consecutive batches of
``--cands`` mixed64 candidates per rank through ``ob2_suggest`` and the sharded
history, with planted duplicates of history rows, of earlier rows of the SAME
step on ANY rank, and of rows of the previous step.  The expected keep mask
does not come from any hash set: a candidate is a function of its Philox
counter and the continuous dimensions never collide otherwise, so

    keep[p] = counter[p] is not a history counter
              and was not drawn in an earlier step
              and p is the first global position (rank * n + i) holding it

is computed from the all-gathered counters with sorts only (the rule of
SURVEY.md Appendix B.3; the reference itself registers blindly,
src/orion/core/worker/producer.py:42-52).  Every rank's mask must equal its
slice bit for bit, and the shards together must have grown by exactly the
number of kept rows.  A last scenario uses a four-row discrete space with one
dominant row, so that one (source, owner) lane overflows its capacity: the ranks must agree to re-run
that batch through worst-case lanes and still produce the exact mask.

Not collected by pytest (needs several GPUs); ``tests/test_gpu_multigpu.py``
launches it when at least two devices are visible.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

HISTORY_BASE = 1 << 40
SEED = 20240005


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--cands', type=int, default=12_500_000)
    ap.add_argument('--history', type=int, default=25_000_000, help='history rows loaded per rank')
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--exchange', default='both', choices=['both', 'peer', 'peer-copy', 'nccl'])
    ap.add_argument('--out', default=None)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    from orion_b200 import spaces
    from orion_b200.dist import PeerShardedHistory, make_sharded_history
    from orion_b200.history import TrialHistory, table_slots_for
    from orion_b200.transformer import build_required_space

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    def total(x):
        t = torch.tensor([int(x)], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(t)
        return int(t.item())

    def gather(t):
        if world == 1:
            return t
        out = torch.empty(world * t.numel(), dtype=t.dtype, device=dev)
        dist.all_gather_into_tensor(out, t.contiguous())
        return out

    n, H, steps = args.cands, args.history, args.steps
    compiled = build_required_space('real', spaces.mixed64()).compiled()
    batch = compiled.empty_columns(n)
    tbatch = compiled.empty_columns(n, transformed=True)
    back = compiled.empty_columns(n)
    keep = torch.empty(batch.n_stride, dtype=torch.uint8, device=dev)

    def step_counters(step, r):
        """Counters of rank r's batch at `step` (any rank can derive any rank's)."""
        g = torch.Generator(device=dev)
        g.manual_seed(SEED + 1000 * step + r)
        gbase = step * world * n
        idx = torch.arange(gbase + r * n, gbase + (r + 1) * n, dtype=torch.int64, device=dev)
        u = torch.rand(n, generator=g, device=dev)
        hist = HISTORY_BASE + torch.randint(0, H * world, (n,), generator=g, device=dev, dtype=torch.int64)
        idx = torch.where(u < 0.10, hist, idx)
        # a copy of one of the first 1 % counters of this step -- whichever rank drew it
        early = gbase + torch.randint(0, max(world * n // 100, 1), (n,), generator=g, device=dev, dtype=torch.int64)
        idx = torch.where((u >= 0.10) & (u < 0.11), early, idx)
        if step > 0:        # a copy of a row of the previous step (history by now, if it was kept then)
            prev = (step - 1) * world * n + torch.randint(0, world * n, (n,), generator=g, device=dev, dtype=torch.int64)
            idx = torch.where((u >= 0.11) & (u < 0.115), prev, idx)
        return idx

    def first_occurrence(values):
        order = torch.argsort(values, stable=True)
        s = values[order]
        first_sorted = torch.ones(values.numel(), dtype=torch.bool, device=dev)
        first_sorted[1:] = s[1:] != s[:-1]
        first = torch.empty_like(first_sorted)
        first[order] = first_sorted
        return first

    report = {"world": world, "cands_per_rank": n, "history_rows_per_rank": H, "steps": steps, "runs": []}
    ok_all = True
    exchanges = ['peer', 'peer-copy', 'nccl'] if args.exchange == 'both' else [args.exchange]
    if world == 1:
        exchanges = ['nccl']            # one rank: no exchange either way
    for exchange in exchanges:
        rows_cap = H + (steps + 1) * n
        history = TrialHistory(compiled, n_slots=table_slots_for(rows_cap, load=0.5))
        sharded = make_sharded_history(history, max_n=n, exchange=exchange)
        used = ('peer-copy' if sharded.copy else 'peer') if isinstance(sharded, PeerShardedHistory) else 'nccl'
        done = 0
        while done < H:
            m = min(n, H - done)
            b = compiled.sample(m, SEED, base_index=HISTORY_BASE + rank * H + done, with_mask=False,
                                with_fingerprint=True, out=batch if m == n else None)
            sharded.insert(b.fp, m)
            done += m
        rows = total(len(history))
        assert rows == H * world, (rows, H * world)
        seen = torch.empty(0, dtype=torch.int64, device=dev)
        for step in range(steps):
            idx = step_counters(step, rank)
            compiled.suggest(n, SEED, index=idx, out=batch, transformed_out=tbatch, back_out=back)
            # alternate the two modes: agreed-on checks per batch / deferred sticky status
            check = step % 2 == 0
            sharded.dedup(back.fp, n, per_rank_n=n, keep=keep, check=check)
            if not check and world > 1:
                sharded.check_overflow() if used != 'nccl' else sharded.check_overflow(n)
            all_idx = gather(idx)
            want = first_occurrence(all_idx) & (all_idx < HISTORY_BASE)
            if seen.numel():
                pos = torch.searchsorted(seen, all_idx).clamp_(max=seen.numel() - 1)
                want &= seen[pos] != all_idx
            mine = want[rank * n:(rank + 1) * n]
            mismatches = total((keep[:n].bool() != mine).sum().item())
            kept = total(keep[:n].sum().item())
            rows_now = total(len(history))
            exact = mismatches == 0 and kept == int(want.sum().item()) and rows_now == rows + kept
            ok_all &= exact
            report["runs"].append({"exchange": used, "step": step, "check_per_batch": check, "mismatches": mismatches,
                                   "kept": kept, "expected_kept": int(want.sum().item()),
                                   "history_rows_before": rows, "history_rows_after": rows_now, "exact": exact})
            rows = rows_now
            seen = torch.unique(torch.cat([seen, all_idx[all_idx < HISTORY_BASE]]))
            del all_idx, want
        del sharded, history
        torch.cuda.empty_cache()

    # ---- skewed owners: four distinct rows, so one lane of the peer exchange overflows
    if world > 1:
        from orion_b200.space import Categorical, Space
        small = Space()
        small.register(Categorical('a', {0: 0.9, 1: 0.1}))      # 81 % of the rows are (0, 0): one lane gets them all
        small.register(Categorical('b', {0: 0.9, 1: 0.1}))
        sc = small.compiled()
        m = 200_000
        for exchange in exchanges:
            history = TrialHistory(sc, capacity=1 << 21)         # (the worst-case lanes of the fallback count as incoming rows)
            sharded = make_sharded_history(history, max_n=m, exchange=exchange)
            used = ('peer-copy' if sharded.copy else 'peer') if isinstance(sharded, PeerShardedHistory) else 'nccl'
            b = sc.sample(m, SEED + 7, base_index=rank * m, with_fingerprint=True)
            k = sharded.dedup(b.fp, m, per_rank_n=m, check=True)
            cols = b.columns
            packed = gather((cols[0][:m].long() * 2 + cols[1][:m].long()).contiguous())
            want = first_occurrence(packed)
            mine = want[rank * m:(rank + 1) * m]
            mismatches = total((k[:m].bool() != mine).sum().item())
            kept = total(k[:m].sum().item())
            rows_now = total(len(history))
            exact = mismatches == 0 and kept == int(want.sum().item()) == rows_now
            ok_all &= exact
            report["runs"].append({"exchange": used, "scenario": "4-row discrete space, overflowing lanes",
                                   "mismatches": mismatches, "kept": kept, "history_rows_after": rows_now,
                                   "exact": exact})
            del sharded, history

    report["keep_mask_exact"] = bool(ok_all)
    if rank == 0:
        text = json.dumps(report)
        print(text)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, 'w') as f:
                f.write(text + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == '__main__':
    main()
