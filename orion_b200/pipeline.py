"""Stream-pipelined candidate generation: the whole hot path, batch after batch.

``fingerprint_in='suggest'`` (what ``bench.py`` times): one kernel per batch
(``ob2_suggest``: sample + contains + transform + reverse + fingerprint, at the
HBM write ceiling) on one stream, dedup (+ exchange) on a second one.

The staged modes are for algorithms that compute in the transformed space
between transform and reverse; a batch then goes through three kernel groups
with different limits on a B200:

  A  sample + transform       (``ob2_sample_transform``)       HBM writes / instruction issue
  B  reverse + fingerprint    (``ob2_reverse_fingerprint``)    HBM bound
  C  dedup (+ exchange)       (``ob2_set_dedup``, NCCL / peer)  random DRAM access

The fingerprint arithmetic costs the sampler 17 % but is free in the
bandwidth-bound reverse, so with ``'reverse'`` B computes it (over the rows it
writes: the candidates as they are handed out) and C follows B.  A of the next
batch depends on neither, so the groups run on three CUDA streams with events
for the true dependencies (and for buffer reuse).  With ``depth`` >= 2 buffer
sets, A of batch k+1 overlaps B and C of batch k -- which buys little: all of
them saturate the same memory system.

This is the producer loop of the reference (``Producer.produce`` ->
``PrimaryAlgo.suggest`` -> ``Random.suggest`` -> ``Space.sample``,
``src/orion/core/worker/producer.py:42-52``) restated for device-resident
batches; every result of the unfused path is still materialised.
"""
import os

import torch

from .dist import ShardedHistory


class BatchSlot(object):
    """One set of buffers travelling through the pipeline."""

    def __init__(self, compiled, n, sample_only=False, alias_back=False):
        dev = compiled.device
        self.batch = compiled.empty_columns(n)                      # original columns + mask + fp + keep
        if sample_only:                                             # no transform chain: the draws are the rows handed out
            self.tbatch, self.back = None, self.batch
        else:
            self.tbatch = compiled.empty_columns(n, transformed=True)   # transform(sample)
            # reverse(transform(sample)): the rows handed out.  alias_back: ONE buffer for them and the draw
            # (ob2_suggest with back == cols) -- the pre-transform draw is an intermediate nobody is handed
            self.back = self.batch if alias_back else compiled.empty_columns(n)
        self.batch.keep = torch.empty(self.batch.n_stride, dtype=torch.uint8, device=dev)
        self.sampled = torch.cuda.Event()
        self.reversed = torch.cuda.Event()
        self.deduped = torch.cuda.Event()
        self.busy = False
        self.hold = None        # event a consumer sets: the slot is not reused before it completes


class CandidatePipeline(object):
    """sample+transform -> (reverse || dedup) over a ring of ``depth`` buffer sets."""

    #: CTAs of the walking dedup grid next to the issue-bound sampler of config 5 (1.5 per SM: the pass then
    #: SHARES the SMs with the next batch's sampler -- step 3.14 -> 2.48 ms -- instead of taking turns with it)
    SAMPLE_ONLY_DEDUP_GRID = 222

    def __init__(self, compiled, history, n, depth=2, timed=False, fingerprint_in='reverse', alias_back=False,
                 dedup_grid=None):
        # 'reverse': the fingerprints are computed by the (bandwidth-bound) reverse kernel over the rows it
        # writes -- the candidates as handed out; 'sample': by the (issue-bound) sampler over its draws
        # 'suggest': ONE kernel (ob2_suggest) samples, transforms, reverses in registers and fingerprints --
        # the Random.suggest fast path; the transformed columns are written but never read back
        # 'sample_only': BASELINE.json config 5 as written -- ob2_sample with the fused fingerprint, then dedup;
        # no transformed columns at all
        assert fingerprint_in in ('reverse', 'sample', 'suggest', 'sample_only')
        self.sample_only = fingerprint_in == 'sample_only'
        self.fused = fingerprint_in in ('suggest', 'sample_only')     # one kernel before the dedup
        self.fp_in_reverse = fingerprint_in in ('reverse', 'suggest')
        self.compiled = compiled
        self.n = n
        # a ShardedHistory / PeerShardedHistory is used as is; a bare TrialHistory is wrapped
        self.sharded = history if hasattr(history, 'world') else ShardedHistory(history, device=compiled.device)
        # next to a kernel that saturates HBM (ob2_suggest's stores, the staged kernels) there is nothing to
        # share: the dedup keeps its one-thread-per-entry grid, the fastest alone
        self.dedup_grid = (self.SAMPLE_ONLY_DEDUP_GRID if self.sample_only else 0) if dedup_grid is None else int(dedup_grid)
        assert not alias_back or fingerprint_in == 'suggest', "only the one-kernel suggest can write both rows into one buffer"
        self.slots = [BatchSlot(compiled, n, sample_only=self.sample_only, alias_back=alias_back) for _ in range(depth)]
        dev = compiled.device
        # The dedup kernels are latency bound (random 32-byte probes) and tiny next to the
        # streaming ones; at equal priority their blocks queue behind the ~24k blocks of
        # sample / reverse and the three groups run back to back.  With the higher priority
        # they are dispatched as soon as SM slots free up and really co-reside.
        self.s_sample = torch.cuda.Stream(device=dev, priority=0)
        self.s_reverse = torch.cuda.Stream(device=dev, priority=0)
        self.s_dedup = torch.cuda.Stream(device=dev, priority=int(os.environ.get('OB2_DEDUP_PRIORITY', '-1')))
        self.k = 0
        self.timed = timed
        self.timings = []          # per step: (sample_ms_events, reverse_events, dedup_events) when timed

    def submit(self, seed, base_index=0, index=None):
        """Enqueue one batch; returns its BatchSlot (results are valid after
        ``slot.reversed`` / ``slot.deduped`` have completed)."""
        slot = self.slots[self.k % len(self.slots)]
        self.k += 1
        c, n = self.compiled, self.n
        caller = torch.cuda.current_stream(c.device)
        ev = None
        if self.timed:
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
        with torch.cuda.stream(self.s_sample):
            self.s_sample.wait_stream(caller)                       # inputs (index) produced by the caller
            if slot.busy:                                           # buffers still read by B / C of an older batch
                self.s_sample.wait_event(slot.reversed)
                self.s_sample.wait_event(slot.deduped)
            if slot.hold is not None:                               # ... or by a consumer (e.g. a copy to the host)
                self.s_sample.wait_event(slot.hold)
                slot.hold = None
            if ev:
                ev[0].record()
            if self.sample_only:
                c.sample(n, seed, base_index=base_index, index=index, with_mask=True, with_fingerprint=True,
                         out=slot.batch)
            elif self.fused:
                c.suggest(n, seed, base_index=base_index, index=index,
                          out=slot.batch, transformed_out=slot.tbatch, back_out=slot.back)
            else:
                c.sample(n, seed, base_index=base_index, index=index, with_mask=True,
                         with_fingerprint=not self.fp_in_reverse,
                         out=slot.batch, transformed_out=slot.tbatch)
            if ev:
                ev[1].record()
            slot.sampled.record()
            if self.fused:
                slot.reversed.record()
        if not self.fused:
            with torch.cuda.stream(self.s_reverse):
                self.s_reverse.wait_event(slot.sampled)
                if ev:
                    ev[2].record()
                c.reverse(slot.tbatch, out=slot.back, with_fingerprint=self.fp_in_reverse)
                if ev:
                    ev[3].record()
                slot.reversed.record()
        with torch.cuda.stream(self.s_dedup):
            self.s_dedup.wait_event(slot.reversed if self.fp_in_reverse else slot.sampled)
            if ev:
                ev[4].record()
            fp = slot.back.fp if self.fp_in_reverse else slot.batch.fp
            self.sharded.set_dedup_grid(self.dedup_grid)            # (a host-side field; the history may be shared)
            self.sharded.dedup(fp, n, per_rank_n=n, keep=slot.batch.keep, check=False)
            if ev:
                ev[5].record()
            slot.deduped.record()
        slot.busy = True
        if ev:
            self.timings.append(ev)
        return slot

    def drain(self):
        """Make the caller's stream wait for everything submitted so far."""
        caller = torch.cuda.current_stream(self.compiled.device)
        caller.wait_stream(self.s_sample)
        caller.wait_stream(self.s_reverse)
        caller.wait_stream(self.s_dedup)

    def kernel_ms(self):
        """Average per-kernel durations (ms) from the on-stream events of timed
        submissions; durations include contention between overlapped kernels."""
        if not self.timings:
            return {}
        torch.cuda.synchronize(self.compiled.device)
        k = len(self.timings)
        if self.fused:
            return {
                ('sample+fingerprint' if self.sample_only else 'suggest'): sum(e[0].elapsed_time(e[1]) for e in self.timings) / k,
                'dedup': sum(e[4].elapsed_time(e[5]) for e in self.timings) / k,
            }
        return {
            'sample+transform': sum(e[0].elapsed_time(e[1]) for e in self.timings) / k,
            'reverse': sum(e[2].elapsed_time(e[3]) for e in self.timings) / k,
            'dedup': sum(e[4].elapsed_time(e[5]) for e in self.timings) / k,
        }
