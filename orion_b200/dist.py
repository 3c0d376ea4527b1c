"""Sharded trial history: dedup across the GPUs of one box.

One process per GPU (``torch.distributed``).  Sampling, bounds checks and the
transform chain shard trivially (rank g owns candidate indices
``[g * n, (g + 1) * n)``) and never communicate.  Dedup has one real exchange:
the history hash set is partitioned by fingerprint (owner = ``fp.hi % world``),
so every rank

  1. buckets its fingerprints by owner (``ob2_bucket_by_owner``),
  2. all-to-alls ``(fp128, u32 global order key)`` to the owners (NCCL),
  3. dedups what it received against its shard, first-wins by global
     candidate index (``ob2_set_dedup`` with the received order keys),
  4. all-to-alls the 1-byte verdicts back, and
  5. scatters them into its local keep mask (``ob2_scatter_verdict``).

Equal rows share a fingerprint, hence an owner, so duplicates that straddle
GPUs resolve exactly as on one GPU.  Buckets have fixed capacity (mean + slack;
fingerprints are uniform) and are zero-padded, so no size pre-exchange is
needed; an overflow is detected from the counts and raised.

The kernel calls are injected (``ops``) so the choreography can be exercised
on CPU with the ``gloo`` backend in the tests.
"""
import ctypes
import math

import torch
import torch.distributed as dist

from . import _lib


def bucket_capacity(n, world):
    """Slots per (source, owner) bucket: mean + 5 % + 6 sigma + 64."""
    mean = n / float(world)
    return int(mean * 1.05 + 6.0 * math.sqrt(max(mean, 1.0))) + 64


class DeviceOps(object):
    """The CUDA kernels behind the exchange (product path)."""

    def __init__(self, device):
        self.lib = _lib.load()
        self.device = torch.device(device)

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def bucket(self, fp, n, key_base, world, cap, out):
        out_fp, out_key, out_src, counts = out
        _lib.check(self.lib.ob2_bucket_by_owner(n, fp.data_ptr(), key_base, world, cap, out_fp.data_ptr(),
                                                out_key.data_ptr(), out_src.data_ptr(), counts.data_ptr(),
                                                self._stream()))

    def dedup(self, history, fp, order, keep):
        history.dedup_fingerprints(fp, fp.shape[0], order=order, keep=keep)

    def insert(self, history, fp):
        history.insert_fingerprints(fp, fp.shape[0])

    def scatter(self, out_src, verdict, out_fp, keep):
        _lib.check(self.lib.ob2_scatter_verdict(out_src.numel(), out_src.data_ptr(), verdict.data_ptr(),
                                                out_fp.data_ptr(), keep.data_ptr(), self._stream()))


class ShardedHistory(object):
    """The history shard of this rank plus the exchange around it."""

    def __init__(self, history, ops=None, group=None, device=None):
        self.history = history
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = torch.device(device) if device is not None else history.device
        self.ops = ops if ops is not None else DeviceOps(self.device)
        self._bufs = {}

    def _buffers(self, n, cap=None):
        cap = bucket_capacity(n, self.world) if cap is None else cap
        key = (n, cap)
        if key not in self._bufs:
            w, dev = self.world, self.device
            self._bufs = {key: dict(
                cap=cap,
                out_fp=torch.empty((w, cap, 2), dtype=torch.int64, device=dev),
                out_key=torch.empty((w, cap), dtype=torch.int32, device=dev),
                out_src=torch.empty((w, cap), dtype=torch.int32, device=dev),
                counts=torch.zeros(w, dtype=torch.int32, device=dev),
                in_fp=torch.empty((w, cap, 2), dtype=torch.int64, device=dev),
                in_key=torch.empty((w, cap), dtype=torch.int32, device=dev),
                verdict=torch.empty(w * cap, dtype=torch.uint8, device=dev),
                back=torch.empty(w * cap, dtype=torch.uint8, device=dev),
            )}
        return self._bufs[key]

    def _route(self, fp, n, per_rank_n, check):
        """Steps 1-2: returns the buffers with ``in_fp`` / ``in_key`` filled.

        With ``check`` the bucket counts are inspected BEFORE anything is sent:
        if any rank has a lane that does not fit (many copies of few rows, e.g.
        a small discrete space -- equal rows share an owner), every rank
        re-buckets into worst-case lanes (capacity n).  Costs one small
        all-reduce and a host sync per batch; ``check=False`` skips it and
        leaves overflow detection to the caller (``_check_overflow``)."""
        b = self._buffers(n)
        key_base = self.rank * per_rank_n
        if (self.world - 1) * per_rank_n + n >= 2 ** 32 - 1:
            raise ValueError("global candidate index does not fit the 32-bit order key")
        self.ops.bucket(fp, n, key_base, self.world, b['cap'], (b['out_fp'], b['out_key'], b['out_src'], b['counts']))
        if check:
            worst = b['counts'].max().to(torch.int64).reshape(1)
            dist.all_reduce(worst, op=dist.ReduceOp.MAX, group=self.group)
            if int(worst.item()) > b['cap']:
                b = self._buffers(n, cap=n)
                self.ops.bucket(fp, n, key_base, self.world, b['cap'],
                                (b['out_fp'], b['out_key'], b['out_src'], b['counts']))
        dist.all_to_all_single(b['in_fp'], b['out_fp'], group=self.group)
        dist.all_to_all_single(b['in_key'], b['out_key'], group=self.group)
        return b

    def _check_overflow(self, b):
        worst = int(b['counts'].max().item())
        if worst > b['cap']:
            raise RuntimeError("exchange bucket overflow: %d fingerprints for one owner, capacity %d"
                               % (worst, b['cap']))

    def dedup(self, fp, n, per_rank_n=None, keep=None, check=True):
        """keep mask (uint8 [>= n]) for this rank's ``n`` fingerprints
        (int64 [m][2], m >= n), deduplicated against the whole sharded history
        and against every rank's batch, first global index wins."""
        per_rank_n = n if per_rank_n is None else per_rank_n
        if keep is None:
            keep = torch.empty(fp.shape[0], dtype=torch.uint8, device=self.device)
        if self.world == 1:
            self.ops.dedup(self.history, fp[:n], None, keep)
            return keep
        b = self._route(fp, n, per_rank_n, check)
        self._owner_side(check, lambda: self.ops.dedup(self.history, b['in_fp'].view(-1, 2), b['in_key'].view(-1),
                                                       b['verdict']))
        dist.all_to_all_single(b['back'], b['verdict'], group=self.group)
        keep.zero_()
        self.ops.scatter(b['out_src'].view(-1), b['back'], b['out_fp'].view(-1, 2), keep)
        return keep

    def insert(self, fp, n, check=True):
        """Add this rank's ``n`` fingerprints to whichever shards own them."""
        if self.world == 1:
            self.ops.insert(self.history, fp[:n])
            return
        b = self._route(fp, n, n, check)
        self._owner_side(check, lambda: self.ops.insert(self.history, b['in_fp'].view(-1, 2)))

    def _owner_side(self, check, call):
        """Run the owner-side set operation.  With ``check`` a failure on ANY rank (a full
        table that may not grow) is raised on EVERY rank -- a rank raising alone would leave
        the others waiting in the next all-to-all."""
        error = None
        try:
            call()
        except _lib.Ob2Error as exc:
            if not check:
                raise
            error = exc
        if check:
            failed = torch.tensor([1 if error else 0], dtype=torch.int32, device=self.device)
            dist.all_reduce(failed, op=dist.ReduceOp.MAX, group=self.group)
            if int(failed.item()):
                raise _lib.Ob2Error("a history shard refused the batch (on this or another rank): %s"
                                    % (error if error else "see the owning rank"))

    def set_dedup_grid(self, ctas):
        self.history.set_dedup_grid(ctas)

    def check_overflow(self, n):
        """After a run of ``check=False`` calls with batches of ``n``: raise if
        the last one dropped fingerprints."""
        if self.world > 1:
            self._check_overflow(self._buffers(n))


class PeerExchange(object):
    """ctypes wrapper of one rank's ``ob2_xchg`` (inbox region + kernels)."""

    def __init__(self, rank, world, cap, device):
        self.lib = _lib.load()
        self.rank, self.world, self.cap = rank, world, cap
        self.device = torch.device(device)
        handle = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.ob2_xchg_create(rank, world, cap, ctypes.byref(handle)))
        self.handle = handle

    def __del__(self):
        handle = getattr(self, 'handle', None)
        if handle:
            try:
                self.lib.ob2_xchg_destroy(handle)
            except Exception:
                pass
            self.handle = None

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def ipc_handle(self):
        buf = ctypes.create_string_buffer(64)
        _lib.check(self.lib.ob2_xchg_ipc_handle(self.handle, buf))
        return buf.raw

    def open_peers(self, handles):
        """handles: list of the 64-byte IPC handles of every rank, in rank order."""
        blob = b''.join(handles)
        assert len(blob) == 64 * self.world
        _lib.check(self.lib.ob2_xchg_open_peers(self.handle, blob))

    def set_peers(self, regions):
        """Same-process ranks: raw region addresses in rank order."""
        arr = (ctypes.c_void_p * self.world)(*regions)
        _lib.check(self.lib.ob2_xchg_set_peers(self.handle, arr))

    def region(self):
        return self.lib.ob2_xchg_region(self.handle)

    def send(self, fp, n, key_base, copy=False):
        fn = self.lib.ob2_xchg_send_copy if copy else self.lib.ob2_xchg_send
        _lib.check(fn(self.handle, n, fp.data_ptr(), key_base, self._stream()))

    def dedup(self, history, insert_only=False, padded=False):
        _lib.check(self.lib.ob2_xchg_dedup(self.handle, history.handle, (1 if insert_only else 0) | (2 if padded else 0),
                                           self._stream()))

    def collect(self, n, keep, copy=False):
        fn = self.lib.ob2_xchg_collect_copy if copy else self.lib.ob2_xchg_collect
        _lib.check(fn(self.handle, n, keep.data_ptr(), self._stream()))

    def send_counts(self):
        out = (ctypes.c_uint32 * self.world)()
        _lib.check(self.lib.ob2_xchg_send_counts(self.handle, out, self._stream()))
        return list(out)


class PeerExchangeUnavailable(RuntimeError):
    """The ranks cannot map each other's inboxes (raised on EVERY rank together)."""


def make_sharded_history(history, max_n, exchange='auto', group=None):
    """The sharded history for this process group: ``'peer'`` (peer-memory
    inboxes filled by remote stores from the bucketing kernel), ``'peer-copy'``
    (the same inboxes filled by DMA transfers), ``'nccl'`` (all-to-all), or
    ``'auto'`` = peer where the ranks can map each other's memory, NCCL
    otherwise.  Collective: call on every rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1 or exchange == 'nccl':
        return ShardedHistory(history, group=group)
    try:
        return PeerShardedHistory(history, max_n, group=group, transfers='copy' if exchange == 'peer-copy' else 'store')
    except PeerExchangeUnavailable:
        if exchange in ('peer', 'peer-copy'):
            raise
        return ShardedHistory(history, group=group)


class PeerShardedHistory(object):
    """Sharded dedup whose exchange runs over peer memory (NVLink stores from
    inside the kernels) instead of NCCL all-to-alls; same results as
    :class:`ShardedHistory`.  Cross-rank ordering uses two small all-reduces
    per batch as stream-ordered barriers.

    Failures are agreed on collectively -- every rank raises, or none does:
    the barrier all-reduce (MAX) carries two sticky status words, *a lane
    overflowed* (many copies of few rows, e.g. a small discrete space: equal
    rows share an owner) and *an owner's table is full*.  With ``check=True``
    the words are read after each phase (one host sync each): an overflowed
    batch is re-run by every rank through the NCCL path with worst-case
    lanes, BEFORE any history has been touched; a full table raises
    everywhere (with ``auto_grow`` the owner doubles its table instead and
    nobody notices).  With ``check=False`` nothing is read on the way and
    :meth:`check_overflow` reports the sticky words after the run."""

    OVERFLOW, FULL = 0, 1

    def __init__(self, history, max_n, group=None, transfers='store'):
        # 'store' (default): the bytes cross NVLink as remote stores / loads issued by the exchange kernels;
        # 'copy': as DMA transfers (ob2_xchg_send_copy / collect_copy).  Measured at 2 GPUs, 1.25e7 per rank:
        # dedup + exchange 1.30 ms with stores, 1.51 with DMA copies (NCCL all-to-all 1.66); step 4.80 / 4.94 / 5.03 ms
        assert transfers in ('copy', 'store')
        self.copy = transfers == 'copy'
        self.history = history
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = history.device
        self.max_n = int(max_n)
        self.xchg = PeerExchange(self.rank, self.world, bucket_capacity(self.max_n, self.world), self.device)
        self.cap = int(self.xchg.lib.ob2_xchg_cap(self.xchg.handle))     # as rounded by the library
        self._status = torch.zeros(2, dtype=torch.int32, device=self.device)
        self._nccl = None
        if self.world > 1:
            mine = torch.tensor(list(self.xchg.ipc_handle()), dtype=torch.uint8, device=self.device)
            everyone = [torch.empty(64, dtype=torch.uint8, device=self.device) for _ in range(self.world)]
            dist.all_gather(everyone, mine, group=group)
            error = None
            try:
                self.xchg.open_peers([bytes(t.cpu().tolist()) for t in everyone])
            except _lib.Ob2Error as exc:       # no peer access / IPC between these processes
                error = exc
            # every rank must reach the same conclusion, or the next collective deadlocks
            ok = torch.tensor([0 if error else 1], dtype=torch.int32, device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0:
                raise PeerExchangeUnavailable(str(error) if error else "a peer rank could not map the inboxes")

    def _barrier(self):
        """Stream-ordered barrier that also spreads the status words (MAX)."""
        if self.world > 1:
            dist.all_reduce(self._status, op=dist.ReduceOp.MAX, group=self.group)

    def _fallback(self):
        if self._nccl is None:
            self._nccl = ShardedHistory(self.history, group=self.group, device=self.device)
        return self._nccl

    def _send(self, fp, n, key_base):
        self.xchg.send(fp, n, key_base, copy=self.copy)
        _lib.check(self.xchg.lib.ob2_xchg_flag_overflow(self.xchg.handle, self._status[self.OVERFLOW:].data_ptr(),
                                                       self.xchg._stream()))
        self._barrier()

    def _owner_side(self, insert_only):
        """Dedup (or load) my inbox.  A full table grows when the history allows
        it; otherwise the inbox is left alone and the FULL word is raised for
        everybody (no exception on this rank alone: the others would hang in
        the next collective)."""
        while True:
            rc = self.xchg.lib.ob2_xchg_dedup(self.xchg.handle, self.history.handle,
                                              (1 if insert_only else 0) | (2 if self.copy else 0), self.xchg._stream())
            if rc == -4 and self.history.auto_grow and self.history.n_slots < (1 << 31):
                self.history.grow()
                continue
            if rc == -4:
                self._full_error = _lib.load().ob2_last_error().decode('utf-8', 'replace')
                self._status[self.FULL:].fill_(1)
            else:
                _lib.check(rc)
            break
        self._barrier()

    def _read_status(self):
        return [int(v) for v in self._status.tolist()]          # host sync; same words on every rank

    def _raise_full(self):
        raise _lib.Ob2Error("a history shard is full (on this or another rank): %s"
                            % getattr(self, '_full_error', 'see the owning rank'))

    def dedup(self, fp, n, per_rank_n=None, keep=None, check=True):
        per_rank_n = n if per_rank_n is None else per_rank_n
        if n > self.max_n:
            raise ValueError("batch of %d exceeds the exchange capacity sized for %d" % (n, self.max_n))
        if (self.world - 1) * per_rank_n + n >= 2 ** 32 - 1:
            raise ValueError("global candidate index does not fit the 32-bit order key")
        if keep is None:
            keep = torch.empty(fp.shape[0], dtype=torch.uint8, device=self.device)
        self._send(fp, n, self.rank * per_rank_n)
        if check and self._read_status()[self.OVERFLOW]:
            # nothing has touched the history yet: every rank takes this batch through worst-case lanes
            self._status[self.OVERFLOW:self.OVERFLOW + 1].zero_()
            return self._fallback().dedup(fp, n, per_rank_n=per_rank_n, keep=keep, check=True)
        self._owner_side(insert_only=False)
        if check and self._read_status()[self.FULL]:
            self._raise_full()
        self.xchg.collect(n, keep, copy=self.copy)
        return keep

    def insert(self, fp, n, check=True):
        if n > self.max_n:
            raise ValueError("batch of %d exceeds the exchange capacity sized for %d" % (n, self.max_n))
        self._send(fp, n, 0)
        if check and self._read_status()[self.OVERFLOW]:
            self._status[self.OVERFLOW:self.OVERFLOW + 1].zero_()
            return self._fallback().insert(fp, n, check=True)
        self._owner_side(insert_only=True)
        if check and self._read_status()[self.FULL]:
            self._raise_full()

    def set_dedup_grid(self, ctas):
        self.history.set_dedup_grid(ctas)

    def check_overflow(self):
        """After ``check=False`` calls: raise -- on every rank alike -- if any
        rank dropped fingerprints or ran out of table since the last check."""
        overflow, full = self._read_status()
        if full:
            self._raise_full()
        if overflow:
            raise RuntimeError("exchange lane overflow on some rank: more than %d fingerprints for one owner; "
                               "verdicts of that batch are incomplete (use check=True)" % self.cap)
