"""Columnar candidate batches: the tensor-returning fast path.

The reference is columnar until the very last line of ``Space.sample``
(``list(zip(*samples))``, space.py:484); a :class:`ColumnBatch` is that pre-zip
form held in device memory: one contiguous block, the column of scalar slot
``s`` starting at byte ``col_units[s] * n_stride``.
"""
import numpy as np
import torch

_TORCH = {'f8': torch.float64, 'i8': torch.int64, 'i4': torch.int32}
_NUMPY = {'f8': np.float64, 'i8': np.int64, 'i4': np.int32}
_SIZE = {'f8': 8, 'i8': 8, 'i4': 4}


class ColumnBatch(object):
    """n candidates (or transformed points) as device columns."""

    def __init__(self, compiled, n, n_stride, buf, transformed=False):
        self.compiled = compiled
        self.n = n
        self.n_stride = n_stride
        self.buf = buf
        self.transformed = transformed
        self.mask = None        # uint8 [n_stride]: point in space (fused contains)
        self.fp = None          # int64 [n_stride][2]: row fingerprints (bit pattern of u64)
        self.keep = None        # uint8 [n_stride]: survived dedup
        self.n_failed = None    # int32 [n_slots]: "Improbable bounds" counters per dimension

    @property
    def _meta(self):
        c = self.compiled
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
    def to_numpy(self):
        """One numpy array per column (a single device->host copy)."""
        host = self.buf.cpu().numpy()
        out = []
        for units, dt in self._meta:
            start = units * self.n_stride
            out.append(host[start:start + _SIZE[dt] * self.n_stride].view(_NUMPY[dt])[:self.n].copy())
        return out

    @classmethod
    def from_numpy(cls, compiled, cols, transformed=False, int_as_f64=False):
        """Upload host columns.  With ``int_as_f64`` Integer columns are sent as
        float64 (raw user points whose integrality the device must check)."""
        n = len(cols[0]) if len(cols) else 0
        batch = compiled.empty_columns(n, transformed)
        meta = batch._meta
        assert len(cols) == len(meta), "expected %d columns, got %d" % (len(meta), len(cols))
        host = np.zeros(batch.buf.numel(), dtype=np.uint8)
        for (units, dt), col in zip(meta, cols):
            want = np.float64 if (int_as_f64 and dt == 'i8' and not transformed) else _NUMPY[dt]
            arr = np.ascontiguousarray(col, dtype=want)
            assert arr.shape == (n,)
            start = units * batch.n_stride
            host[start:start + arr.nbytes] = arr.view(np.uint8)
        batch.buf.copy_(torch.from_numpy(host))
        return batch

    def compact(self, keep=None, out=None, count=None, scratch=None):
        """Dense ColumnBatch of the rows with ``keep != 0`` (default: the
        batch's own dedup verdicts), original order preserved
        (``ob2_compact``).  Reads the survivor count back (one sync) unless a
        device ``count`` tensor (int64[1]) is passed, in which case nothing
        synchronises and ``dest.n`` is left for the caller to set."""
        import ctypes
        from . import _lib
        c = self.compiled
        keep = self.keep if keep is None else keep
        assert keep is not None and keep.dtype == torch.uint8
        dest = out if out is not None else c.empty_columns(self.n, self.transformed)
        nbytes = c.lib.ob2_compact_scratch_bytes(self.n)
        if scratch is None:
            scratch = torch.empty(max(nbytes, 8), dtype=torch.uint8, device=c.device)
        assert scratch.numel() * scratch.element_size() >= nbytes
        lazy = count is not None
        if count is None:
            count = torch.zeros(1, dtype=torch.int64, device=c.device)
        _lib.check(c.lib.ob2_compact(c.handle, self.n, self.n_stride, self.buf.data_ptr(), keep.data_ptr(),
                                     1 if self.transformed else 0, dest.n_stride, dest.buf.data_ptr(),
                                     scratch.data_ptr(), count.data_ptr(),
                                     ctypes.c_void_p(torch.cuda.current_stream(c.device).cuda_stream)))
        if not lazy:
            dest.n = int(count.item())
        return dest

    def fingerprints_numpy(self):
        """(lo, hi) uint64 arrays of the first n fingerprints."""
        f = self.fp[:self.n].cpu().numpy().view(np.uint64)
        return f[:, 0].copy(), f[:, 1].copy()
