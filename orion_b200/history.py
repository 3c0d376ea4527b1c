"""Trial-history dedup on the device (single GPU).

The reference snapshot registers every suggestion blindly
(``src/orion/core/worker/producer.py:42-52``; ``tests/unittests/core/test_producer.py:69-88``
shows two identical points both stored).  The contract implemented here is
SURVEY.md Appendix B.3: a candidate is kept iff its row is not in the history
and no earlier row of the same batch equals it; kept rows join the history.
Rows are identified by a 128-bit fingerprint of their canonical bytes
(collision probability at 1e9 rows ~ 2e-21).
"""
import ctypes

import numpy
import torch

from . import _lib
from .columns import ColumnBatch


def table_slots_for(capacity, load=0.5):
    """Power-of-two slot count that holds ``capacity`` rows at <= ``load``."""
    want = max(64, int(capacity / load))
    return 1 << (want - 1).bit_length()


class TrialHistory(object):
    """Hash set of row fingerprints in HBM + the dedup operations on it."""

    def __init__(self, space, capacity=1 << 20, n_slots=None, device=None, auto_grow=False):
        self.lib = _lib.load()
        self.space = space
        self.compiled = space.compiled() if hasattr(space, 'compiled') else space
        self.device = self.compiled.device if device is None else torch.device(device)
        self.auto_grow = auto_grow
        self.n_slots = int(n_slots) if n_slots else table_slots_for(capacity)
        self.table, self.handle = self._new_table(self.n_slots)
        self._scratch = None

    def _new_table(self, n_slots):
        nbytes = _lib.SET_HEADER_BYTES + n_slots * _lib.SET_SLOT_BYTES
        table = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        assert table.data_ptr() % 256 == 0
        handle = ctypes.c_void_p()
        _lib.check(self.lib.ob2_set_create(table.data_ptr(), n_slots, self._stream(), ctypes.byref(handle)))
        return table, handle

    def set_dedup_grid(self, ctas):
        """0: the dedup's first pass runs one thread per entry (fastest alone); ctas > 0: that many CTAs
        walk the entries, so the pass shares the SMs with a kernel on another stream instead of taking
        turns with it (``ob2_set_set_dedup_grid``; what the config-5 pipeline uses next to its sampler)."""
        _lib.check(self.lib.ob2_set_set_dedup_grid(self.handle, int(ctas)))
        self._dedup_grid = int(ctas)

    def grow(self, n_slots=None):
        """Move the history into a table of ``n_slots`` (default: twice as many);
        everything stored so far stays a duplicate for future batches."""
        n_slots = int(n_slots) if n_slots else 2 * self.n_slots
        if n_slots <= self.n_slots or n_slots & (n_slots - 1):
            raise ValueError("grow() needs a larger power-of-two slot count")
        table, handle = self._new_table(n_slots)
        _lib.check(self.lib.ob2_set_rehash(self.handle, handle, self._stream()))
        self.lib.ob2_set_destroy(self.handle)
        self.table, self.handle, self.n_slots = table, handle, n_slots
        if getattr(self, '_dedup_grid', None) is not None:          # the setting belongs to the history, not to one table
            self.set_dedup_grid(self._dedup_grid)

    def _guarded(self, call):
        """Run a set operation; with ``auto_grow`` a full table is doubled and
        the operation retried (the library refuses BEFORE touching the table)."""
        while True:
            rc = call()
            if rc == -4 and self.auto_grow and self.n_slots < (1 << 31):      # OB2_ENOMEM: set full
                self.grow()
                continue
            _lib.check(rc)
            return

    # -- snapshot / restore ----------------------------------------------------
    def state_dict(self):
        """The stored fingerprints as a host tensor (int64 [count][2]): enough
        to rebuild the history in a table of any size."""
        count = len(self)
        out = torch.empty((max(count, 1), 2), dtype=torch.int64, device=self.device)
        n_out = torch.zeros(1, dtype=torch.int64, device=self.device)
        _lib.check(self.lib.ob2_set_export(self.handle, count, out.data_ptr(), n_out.data_ptr(), self._stream()))
        assert int(n_out.item()) == count
        return {'fingerprints': out[:count].cpu(), 'n_slots': self.n_slots}

    def load_state_dict(self, state):
        fp = state['fingerprints'].to(self.device)
        if fp.shape[0]:
            self.insert_fingerprints(fp.contiguous())

    def __del__(self):
        handle = getattr(self, 'handle', None)
        if handle:
            try:
                self.lib.ob2_set_destroy(handle)
            except Exception:
                pass
            self.handle = None

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _scratch_for(self, n):
        if self._scratch is None or self._scratch.numel() < n:
            self._scratch = torch.empty(max(n, 1024), dtype=torch.int32, device=self.device)
        return self._scratch

    # -- fingerprints in, verdicts out (device tensors) ----------------------
    def insert_fingerprints(self, fp, n=None):
        """History load: ``fp`` is an int64 [m][2] device tensor."""
        n = fp.shape[0] if n is None else n
        self._guarded(lambda: self.lib.ob2_set_insert(self.handle, n, fp.data_ptr(), self._stream()))

    def dedup_fingerprints(self, fp, n=None, order=None, keep=None):
        """uint8 keep mask for ``fp`` (first occurrence by ``order`` wins)."""
        n = fp.shape[0] if n is None else n
        if keep is None:
            keep = torch.empty(fp.shape[0], dtype=torch.uint8, device=self.device)
        scratch = self._scratch_for(n)
        self._guarded(lambda: self.lib.ob2_set_dedup(self.handle, n, fp.data_ptr(),
                                                     order.data_ptr() if order is not None else None,
                                                     scratch.data_ptr(), keep.data_ptr(), self._stream()))
        return keep

    def dedup_fingerprints_static(self, fp, keep, n=None, order=None):
        """``dedup_fingerprints`` for callers that may not allocate on the
        device: the per-entry scratch is the fixed array built into libob2
        (``ob2_set_dedup_static``), so the caller passes only ``fp`` in and
        ``keep`` out, both already allocated.  At most
        ``_lib.STATIC_DEDUP_MAX`` entries per call; calls must not overlap on
        the GPU (they share that one array); make the first call on a device
        before any stream capture (it looks the array's address up)."""
        n = fp.shape[0] if n is None else n
        _lib.check(self.lib.ob2_set_dedup_static(self.handle, n, fp.data_ptr(),
                                                 order.data_ptr() if order is not None else None,
                                                 keep.data_ptr(), self._stream()))
        return keep

    def probe_fingerprints(self, fp, n=None):
        n = fp.shape[0] if n is None else n
        found = torch.empty(fp.shape[0], dtype=torch.uint8, device=self.device)
        _lib.check(self.lib.ob2_set_probe(self.handle, n, fp.data_ptr(), found.data_ptr(), self._stream()))
        return found

    # -- batches -------------------------------------------------------------
    def observe(self, batch):
        """Record every row of a ColumnBatch (e.g. completed trials)."""
        if batch.fp is None:
            self.compiled.fingerprint(batch)
        self.insert_fingerprints(batch.fp, batch.n)

    def dedup(self, batch):
        """Set ``batch.keep`` (uint8 [n_stride]); kept rows join the history."""
        if batch.fp is None:
            self.compiled.fingerprint(batch)
        batch.keep = self.dedup_fingerprints(batch.fp, batch.n, keep=batch.keep)
        return batch.keep[:batch.n]

    # -- points (compatibility path) -----------------------------------------
    def _batch_from_points(self, points):
        cols, row_ok = self.space.columns_from_rows(points)
        if not row_ok.all():
            raise ValueError("point does not match the shape of the space")
        typed = [numpy.asarray(c, dtype={'f8': numpy.float64, 'i8': numpy.int64, 'i4': numpy.int32}[dt])
                 for c, (_u, dt) in zip(cols, self.compiled.col_meta)]
        return ColumnBatch.from_numpy(self.compiled, typed)

    def observe_points(self, points):
        self.observe(self._batch_from_points(points))

    def dedup_points(self, points):
        """List of tuples -> list of bool (keep)."""
        keep = self.dedup(self._batch_from_points(points))
        return list(keep.cpu().numpy().astype(bool))

    def __len__(self):
        count = ctypes.c_uint64()
        _lib.check(self.lib.ob2_set_count(self.handle, ctypes.byref(count), self._stream()))
        return int(count.value)
