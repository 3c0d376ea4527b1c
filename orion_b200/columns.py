"""Columnar candidate batches: the tensor-returning fast path.

The reference is columnar until the very last line of ``Space.sample``
(``list(zip(*samples))``, space.py:484); a :class:`ColumnBatch` is that pre-zip
form held in device memory: one contiguous block, the column of scalar slot
``s`` starting at byte ``col_units[s] * n_stride``.
"""
import numpy as np
import torch

_TORCH = {'f8': torch.float64, 'i8': torch.int64, 'i4': torch.int32, 'i2': torch.int16, 'i1': torch.int8,
          'u2': torch.uint16, 'u1': torch.uint8}
_NUMPY = {'f8': np.float64, 'i8': np.int64, 'i4': np.int32, 'i2': np.int16, 'i1': np.int8,
          'u2': np.uint16, 'u1': np.uint8}
_SIZE = {'f8': 8, 'i8': 8, 'i4': 4, 'i2': 2, 'i1': 1, 'u2': 2, 'u1': 1}
_PINNED_FROM = 4 << 20          # host round trips of at least this many bytes go through pinned memory ...
_PINNED_UP_TO = 1 << 30         # ... unless they would pin more than this at once


class ColumnBatch(object):
    """n candidates (or transformed points) as device columns."""

    def __init__(self, compiled, n, n_stride, buf, transformed=False, wire=False):
        self.compiled = compiled
        self.n = n
        self.n_stride = n_stride
        self.buf = buf
        self.transformed = transformed
        self.wire = wire        # narrowed hand-off layout: same values, narrowest lossless integer widths
        self.mask = None        # uint8 [n_stride]: point in space (fused contains)
        self.fp = None          # int64 [n_stride][2]: row fingerprints (bit pattern of u64)
        self.keep = None        # uint8 [n_stride]: survived dedup
        self.n_failed = None    # int32 [n_slots]: "Improbable bounds" counters per dimension

    @property
    def _meta(self):
        c = self.compiled
        if self.wire:
            return c.wire_meta
        return [(u, dt) for (u, dt, _s) in c.tcol_meta] if self.transformed else c.col_meta

    def column(self, j):
        """Column j as a typed device tensor view of length n."""
        units, dt = self._meta[j]
        start = units * self.n_stride
        raw = self.buf[start:start + _SIZE[dt] * self.n_stride]
        return raw.view(_TORCH[dt])[:self.n]

    @property
    def columns(self):
        return [self.column(j) for j in range(len(self._meta))]

    def __len__(self):
        return self.n

    # ---- host round trips (compatibility path and tests)
    def to_numpy(self, natural=True):
        """One numpy array per column (a single device->host copy of the first n rows).  A wire
        batch is widened on the host to the natural dtypes (int64 / int32) unless ``natural`` is
        False.  Big batches land in ONE pinned block (``ob2_columns_to_host``) that the returned
        arrays are views of -- no pageable bounce, no second copy on the host."""
        meta = self._meta
        row_units = max(u + _SIZE[dt] for u, dt in meta)
        if _PINNED_FROM <= self.n * row_units <= _PINNED_UP_TO:
            hs = -(-self.n // 32) * 32
            pinned = torch.empty(row_units * hs, dtype=torch.uint8, pin_memory=True)
            self.copy_rows_to_host(pinned, self.n, hs)
            torch.cuda.current_stream(self.buf.device).synchronize()
            host = pinned.numpy()
            out = []
            for j, (units, dt) in enumerate(meta):
                col = host[units * hs:units * hs + _SIZE[dt] * self.n].view(_NUMPY[dt])
                out.append(col.astype(_NUMPY[self.compiled.col_meta[j][1]]) if self.wire and natural else col)
            return out
        sparse = 2 * self.n < self.n_stride         # e.g. the survivors of a compacted batch: rows [0, n) only
        if sparse:
            # gather the used prefix of every column into one dense device block, then ONE copy of
            # just those bytes (not row_bytes * n_stride of mostly unused slots)
            sizes = [_SIZE[dt] * self.n for _u, dt in meta]
            offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
            dense = torch.empty(int(offs[-1]), dtype=torch.uint8, device=self.buf.device)
            for (units, _dt), o, sz in zip(meta, offs[:-1], sizes):
                start = units * self.n_stride
                dense[int(o):int(o) + sz].copy_(self.buf[start:start + sz])
            host = dense.cpu().numpy()
        else:
            host = self.buf.cpu().numpy()
        out = []
        for j, (units, dt) in enumerate(meta):
            if sparse:
                col = host[int(offs[j]):int(offs[j + 1])].view(_NUMPY[dt])
            else:
                start = units * self.n_stride
                col = host[start:start + _SIZE[dt] * self.n_stride].view(_NUMPY[dt])[:self.n]
            if self.wire and natural:
                out.append(col.astype(_NUMPY[self.compiled.col_meta[j][1]]))
            else:
                out.append(col.copy())
        return out

    @classmethod
    def from_numpy(cls, compiled, cols, transformed=False, int_as_f64=False):
        """Upload host columns.  With ``int_as_f64`` Integer columns are sent as
        float64 (raw user points whose integrality the device must check)."""
        n = len(cols[0]) if len(cols) else 0
        batch = compiled.empty_columns(n, transformed)
        meta = batch._meta
        assert len(cols) == len(meta), "expected %d columns, got %d" % (len(meta), len(cols))
        big = _PINNED_FROM <= batch.buf.numel() <= _PINNED_UP_TO     # a pinned block and an asynchronous copy; only the tails are zeroed
        staged = torch.empty(batch.buf.numel(), dtype=torch.uint8, pin_memory=True) if big else None
        host = staged.numpy() if big else np.zeros(batch.buf.numel(), dtype=np.uint8)
        for (units, dt), col in zip(meta, cols):
            want = np.float64 if (int_as_f64 and dt == 'i8' and not transformed) else _NUMPY[dt]
            arr = np.ascontiguousarray(col, dtype=want)
            assert arr.shape == (n,)
            start = units * batch.n_stride
            host[start:start + arr.nbytes] = arr.view(np.uint8)
            if big:
                host[start + arr.nbytes:start + _SIZE[dt] * batch.n_stride] = 0
        batch.buf.copy_(staged if big else torch.from_numpy(host), non_blocking=big)
        return batch

    def compact(self, keep=None, out=None, count=None, scratch=None, wire=False):
        """Dense ColumnBatch of the rows with ``keep != 0`` (default: the
        batch's own dedup verdicts), original order preserved
        (``ob2_compact``).  Reads the survivor count back (one sync) unless a
        device ``count`` tensor (int64[1]) is passed, in which case nothing
        synchronises and ``dest.n`` is left for the caller to set.  ``wire``:
        the result is in the narrowed hand-off layout (``ob2_compact_wire``;
        original column sets only) -- fewer bytes to copy to the host."""
        import ctypes
        from . import _lib
        c = self.compiled
        keep = self.keep if keep is None else keep
        assert keep is not None and keep.dtype == torch.uint8
        assert not self.wire and not (wire and self.transformed)
        dest = out if out is not None else c.empty_columns(self.n, self.transformed, wire=wire)
        assert dest.wire == wire and dest.transformed == self.transformed
        assert dest.n_stride >= self.n, "compact(out=...): the destination holds %d rows, up to %d may survive" % (
            dest.n_stride, self.n)
        nbytes = c.lib.ob2_compact_scratch_bytes(self.n)
        if scratch is None:
            scratch = torch.empty(max(nbytes, 8), dtype=torch.uint8, device=c.device)
        assert scratch.numel() * scratch.element_size() >= nbytes
        lazy = count is not None
        if count is None:
            count = torch.zeros(1, dtype=torch.int64, device=c.device)
        stream = ctypes.c_void_p(torch.cuda.current_stream(c.device).cuda_stream)
        if wire:
            _lib.check(c.lib.ob2_compact_wire(c.handle, self.n, self.n_stride, self.buf.data_ptr(), keep.data_ptr(),
                                              dest.n_stride, dest.buf.data_ptr(), scratch.data_ptr(),
                                              count.data_ptr(), stream))
        else:
            _lib.check(c.lib.ob2_compact(c.handle, self.n, self.n_stride, self.buf.data_ptr(), keep.data_ptr(),
                                         1 if self.transformed else 0, dest.n_stride, dest.buf.data_ptr(),
                                         scratch.data_ptr(), count.data_ptr(), stream))
        if not lazy:
            dest.n = int(count.item())
        return dest

    def copy_rows_to_host(self, host_buf, n_rows=None, host_stride=None):
        """Asynchronously copy the first ``n_rows`` of every column into ``host_buf`` (a pinned
        uint8 tensor laid out like this batch with stride ``host_stride``, default the batch's):
        ``ob2_columns_to_host`` -- 2-3 two-dimensional copies, whatever the number of columns."""
        import ctypes
        from . import _lib
        c = self.compiled
        n_rows = self.n if n_rows is None else int(n_rows)
        host_stride = self.n_stride if host_stride is None else int(host_stride)
        layout = 2 if self.wire else (1 if self.transformed else 0)
        stream = ctypes.c_void_p(torch.cuda.current_stream(c.device).cuda_stream)
        _lib.check(c.lib.ob2_columns_to_host(c.handle, layout, n_rows, self.n_stride, self.buf.data_ptr(),
                                             host_stride, host_buf.data_ptr(), stream))
        return host_buf

    def fingerprints_numpy(self):
        """(lo, hi) uint64 arrays of the first n fingerprints."""
        f = self.fp[:self.n].cpu().numpy().view(np.uint64)
        return f[:, 0].copy(), f[:, 1].copy()
