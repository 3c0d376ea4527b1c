"""Space -> device descriptor compiler and the batch-level device operations.

`compile_space` accepts any object that quacks like the reference's
``orion.algo.space.Space`` (an ordered mapping of dimensions exposing ``type``,
``prior``, ``_prior_name``, ``_args``, ``_kwargs``, ``shape``, ``interval()``,
and for categoricals ``categories`` / ``_probs``) -- i.e. both the reference's
own classes and the mirror in :mod:`orion_b200.space`.

Everything static about a dimension is resolved here, once, on the host:
``interval()`` (which the reference recomputes through two scipy ``ppf`` calls
for every scalar it checks, ``space.py:160``), ``log(a)`` / ``log(b) - log(a)``
of a loguniform prior, and the categorical cdf exactly as numpy's legacy
``RandomState.choice`` builds it.  The kernels then only do arithmetic.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import SlotDesc

REQUIREMENTS = ('real', 'integer', 'numerical', None)

_KIND = {'real': _lib.KIND_REAL, 'integer': _lib.KIND_INTEGER, 'categorical': _lib.KIND_CATEGORICAL}
_OPS = {
    'real': {'real': _lib.OP_IDENTITY, 'integer': _lib.OP_TO_REAL, 'categorical': _lib.OP_ONEHOT},
    'integer': {'real': _lib.OP_QUANTIZE, 'integer': _lib.OP_IDENTITY, 'categorical': _lib.OP_ENUMERATE},
    'numerical': {'real': _lib.OP_IDENTITY, 'integer': _lib.OP_IDENTITY, 'categorical': _lib.OP_ONEHOT},
    None: {'real': _lib.OP_IDENTITY, 'integer': _lib.OP_IDENTITY, 'categorical': _lib.OP_IDENTITY},
}

_M64 = (1 << 64) - 1
_ELEM = {'f8': 8, 'i8': 8, 'i4': 4}


def _splitmix64(x):
    x = (x + 0x9E3779B97F4A7C15) & _M64
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def _hash_keys(slot):
    """Per-slot fingerprint constants: [xor key, lane-a multipliers, lane-b
    multipliers, 0]; a multiplier word packs two odd 32-bit multipliers (for
    the low and the high half of the value)."""
    k = [_splitmix64((4 * slot + j + 1) * 0xD6E8FEB86659FD93 & _M64) for j in range(3)]
    return [k[0], k[1] | 0x0000000100000001, k[2] | 0x0000000100000001, 0]


def round_up(n, m=32):
    return (int(n) + m - 1) // m * m


# scipy priors with a closed-form inverse CDF on the device:
# name -> (dist code, number of shape parameters, sign of the reciprocal stored in p1 or 0)
_CLOSED_FORM = {
    'expon': (5, 0, 0), 'lognorm': (6, 1, 0), 'logistic': (7, 0, 0), 'laplace': (8, 0, 0), 'gumbel_r': (9, 0, 0),
    'weibull_min': (10, 1, 1.0), 'pareto': (11, 1, -1.0), 'rayleigh': (12, 0, 0), 'halfnorm': (13, 0, 0),
    'powerlaw': (14, 1, 1.0),
}


# more closed forms (ob2.h OB2_DIST_GUMBEL_L..): name -> (dist code, number of shape parameters); p0 / p1 carry them as given
_CLOSED_FORM_2 = {
    'gumbel_l': (26, 0), 'genpareto': (27, 1), 'lomax': (28, 1), 'burr': (29, 2), 'burr12': (30, 2), 'invweibull': (31, 1),
    'truncexpon': (32, 1), 'bradford': (33, 1), 'arcsine': (34, 0), 'triang': (35, 1), 'loglaplace': (36, 1),
    'exponpow': (37, 1), 'gompertz': (38, 1), 'genlogistic': (39, 1), 'hypsecant': (40, 0), 'halfcauchy': (41, 0),
    'halflogistic': (42, 0), 'gibrat': (46, 0), 'genextreme': (47, 1), 'kappa3': (48, 1), 'dweibull': (49, 1),
    'anglit': (50, 0), 'exponweib': (51, 2), 'mielke': (52, 2), 'genhalflogistic': (53, 1), 'tukeylambda': (54, 1),
    'levy': (55, 0), 'levy_l': (56, 0), 'weibull_max': (57, 1), 'moyal': (61, 0), 'wrapcauchy': (62, 1),
    'johnsonsb': (68, 2), 'johnsonsu': (69, 2), 'fatiguelife': (70, 1), 'powernorm': (71, 1), 'powerlognorm': (72, 2),
    'kappa4': (74, 2), 'laplace_asymmetric': (75, 1), 'trapezoid': (76, 2), 'truncpareto': (77, 2), 'truncweibull_min': (78, 3),
    'skewcauchy': (79, 1), 'planck': (80, 1), 'dlaplace': (81, 1), 'boltzmann': (82, 2),
}
# ... and those that go through the incomplete gamma inverse: name -> (dist code, a as a function of the shape parameters)
_GAMMA_BASED = {
    'chi': (43, lambda df: 0.5 * df), 'maxwell': (43, lambda: 1.5), 'invgamma': (44, lambda a: a),
    'nakagami': (45, lambda nu: nu), 'dgamma': (59, lambda a: a), 'gennorm': (60, lambda beta: 1.0 / beta),
}


def _draws_by_inverse_cdf(prior):
    """True when scipy's rvs for this prior is the generic ``_ppf(random_state.uniform(size))``
    (rv_generic._rvs) rather than a generator of its own."""
    from scipy.stats._distn_infrastructure import rv_continuous, rv_generic
    return isinstance(prior, rv_continuous) and getattr(type(prior), '_rvs', None) is rv_generic._rvs


def _prior_params(dim):
    """(dist code, p0, p1, loc, scale) of a Real/Integer dimension."""
    name = dim._prior_name
    try:
        args, loc, scale, _ = dim.prior._parse_args_rvs(*dim._args, size=None, **dim._kwargs)
    except TypeError as exc:
        raise TypeError("Parameter '{}': incorrect arguments for prior '{}': {}".format(dim.name, name, exc))
    loc, scale = float(loc), float(scale)
    if not (scale >= 0) or not np.all(dim.prior._argcheck(*args)):
        raise ValueError("Domain error in arguments of prior '{}' of '{}'".format(name, dim.name))
    if scale == 0:
        return _lib.DIST_CONST, 0.0, 0.0, loc, scale
    if name == 'uniform':
        return _lib.DIST_UNIFORM, 0.0, 0.0, loc, scale
    if name in ('reciprocal', 'loguniform'):
        a, b = float(args[0]), float(args[1])
        la = np.log(a)
        return _lib.DIST_RECIPROCAL, float(la), float(np.log(b) - la), loc, scale
    if name == 'norm':
        return _lib.DIST_NORM, 0.0, 0.0, loc, scale
    if name in _CLOSED_FORM:
        code, n_shape, reciprocal_sign = _CLOSED_FORM[name]
        if len(args) != n_shape:
            raise TypeError("Parameter '{}': prior '{}' takes {} shape parameter(s)".format(dim.name, name, n_shape))
        p0 = float(args[0]) if n_shape else 0.0
        p1 = reciprocal_sign / p0 if (n_shape and reciprocal_sign) else 0.0
        return code, p0, p1, loc, scale
    if name in _CLOSED_FORM_2 and len(args) == _CLOSED_FORM_2[name][1]:
        shapes = [float(a) for a in args] + [0.0, 0.0, 0.0]
        return _CLOSED_FORM_2[name][0], shapes[0], shapes[1], loc, scale, shapes[2]
    if name == 'gengamma' and len(args) == 2 and float(args[1]) != 0.0:
        from scipy.special import gammaln
        a = float(args[0])
        return 73, a, float(gammaln(a)), loc, scale, float(args[1])
    if name == 'bernoulli' and len(args) == 1:          # binom(1, p)
        p = float(args[0])
        if p < 1:
            return 25, 1.0, p, loc, scale, 1.0 - p
    if name == 'fisk' and len(args) == 1:               # burr(c, 1)
        return 29, float(args[0]), 1.0, loc, scale
    if name in _GAMMA_BASED:
        from scipy.special import gammaln
        code, shape_of = _GAMMA_BASED[name]
        a = float(shape_of(*[float(x) for x in args]))
        return code, a, float(gammaln(a)), loc, scale, (float(args[0]) if name == 'gennorm' else 0.0)
    if name in ('rdist', 'semicircular') and len(args) == (1 if name == 'rdist' else 0):
        from scipy.special import betaln
        half = 0.5 * (float(args[0]) if name == 'rdist' else 3.0)
        return 58, half, 0.0, loc, scale, float(betaln(half, half))
    if name == 'f' and len(args) == 2:
        from scipy.special import betaln
        dfn, dfd = float(args[0]), float(args[1])
        return 63, dfn, dfd, loc, scale, float(betaln(0.5 * dfn, 0.5 * dfd))
    if name == 'loggamma' and len(args) == 1:
        from scipy.special import gammaln
        c = float(args[0])
        return 64, c, float(gammaln(c)), loc, scale, float(gammaln(c + 1.0))
    if name == 'nbinom' and len(args) == 2:
        n_succ, p = float(args[0]), float(args[1])
        pmf0 = float(np.exp(n_succ * np.log(p))) if p > 0 else 0.0
        if pmf0 > 1e-290:          # (otherwise the pmf recurrence cannot start: bounds-only)
            return 65, n_succ, p, loc, scale, pmf0
    if name == 'hypergeom' and len(args) == 3:
        return 66, float(args[0]), float(args[1]), loc, scale, float(args[2])
    if name == 'vonmises' and len(args) == 1 and float(args[0]) < 50.0:     # (scipy switches to a normal approximation at 50)
        from scipy.special import i0e
        kappa = float(args[0])
        return 67, kappa, float(1.0 / (2.0 * np.pi * i0e(kappa))), loc, scale
    if name == 'alpha' and len(args) == 1:
        from scipy.special import ndtr
        a = float(args[0])
        return 15, a, float(ndtr(a)), loc, scale
    if name == 'randint' and len(args) == 2:
        low, high = float(args[0]), float(args[1])
        return 16, low, high - low, loc, scale
    if name == 'poisson' and len(args) == 1 and float(args[0]) <= 500.0:
        mu = float(args[0])
        return 17, mu, float(np.exp(-mu)), loc, scale
    if name in ('gamma', 'erlang') and len(args) == 1:
        from scipy.special import gammaln
        a = float(args[0])
        return 18, a, float(gammaln(a)), loc, scale
    if name == 'beta' and len(args) == 2:
        from scipy.special import betaln
        a, b = float(args[0]), float(args[1])
        return 19, a, b, loc, scale, float(betaln(a, b))
    if name == 'cauchy' and not args:
        return 20, 0.0, 0.0, loc, scale
    if name == 'chi2' and len(args) == 1:
        from scipy.special import gammaln
        half = 0.5 * float(args[0])
        return 21, half, float(gammaln(half)), loc, scale
    if name == 't' and len(args) == 1:
        from scipy.special import betaln
        df = float(args[0])
        return 22, df, float(betaln(0.5 * df, 0.5)), loc, scale
    if name == 'truncnorm' and len(args) == 2:
        from scipy.special import ndtr
        from scipy.stats._continuous_distns import _log_gauss_mass
        a, b = float(args[0]), float(args[1])
        mass = float(np.exp(_log_gauss_mass(np.float64(a), np.float64(b))))
        if a < 0:           # scipy's left-hand case: Phi(x) = Phi(a) + q mass
            return 23, float(ndtr(a)), mass, loc, scale, 1.0
        return 23, float(ndtr(-b)), mass, loc, scale, -1.0      # right-hand: Phi(-x) = Phi(-b) + (1 - q) mass
    if name == 'geom' and len(args) == 1:
        p = float(args[0])
        return 24, p, float(np.log1p(-p)), loc, scale
    if name == 'binom' and len(args) == 2:
        n_trials, p = float(args[0]), float(args[1])
        pmf0 = float(np.exp(n_trials * np.log1p(-p))) if p < 1 else 0.0
        if pmf0 > 1e-290:          # (otherwise the pmf recurrence cannot start: bounds-only)
            return 25, n_trials, p, loc, scale, pmf0
    # any other scipy prior: bounds-only (contains / transform / dedup work, sampling refuses)
    return _lib.DIST_NONE, 0.0, 0.0, loc, scale


class CompiledSpace(object):
    """A Space compiled for the device, under one transform requirement."""

    def __init__(self, space, requirement=None, quantize='round', device=None, philox_rounds=10):
        if requirement not in REQUIREMENTS:
            raise TypeError("Unsupported requirement: {}".format(requirement))
        if quantize not in ('round', 'floor'):
            raise ValueError("quantize must be 'round' or 'floor'")
        self.lib = _lib.load()
        self.requirement = requirement
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        dims = list(space.values())
        if not dims:
            raise ValueError("cannot compile an empty Space")
        self.dim_names = [d.name for d in dims]
        self.dim_kinds = ['real' if d.type == 'dimension' else d.type for d in dims]
        self.dim_shapes = [tuple(d.shape) for d in dims]
        self.dim_categories = [tuple(d.categories) if d.type == 'categorical' else None for d in dims]
        self.unsamplable = None        # (dimension name, prior) without a device inverse CDF
        self.dim_first = []            # first slot of each dimension
        self.dim_bounds = []
        self.dim_dists = []            # prior code of each dimension (None: categorical)
        self.dim_generic_rvs = []      # does scipy draw this prior as _ppf(random_state.uniform())? (mt_stream)
        slots = []
        cdf = []
        col_units = t_units = 0
        self.col_meta = []             # per slot: (byte offset units, numpy dtype str)
        self.tcol_meta = []            # per transformed column
        for d_idx, dim in enumerate(dims):
            kind = dim.type
            if kind == 'dimension':     # bare `Dimension`: a Real without extra bounds
                kind = 'real'
            if kind not in _KIND:
                raise TypeError("Dimension '{}' of type '{}' cannot be compiled".format(dim.name, kind))
            k = int(np.prod(dim.shape, dtype=np.int64)) if len(dim.shape) else 1
            first = len(slots)
            self.dim_first.append(first)
            if kind == 'categorical':
                probs = np.asarray(dim._probs, dtype=np.float64)
                c = probs.cumsum()
                c /= c[-1]
                cdf_off = len(cdf)
                cdf.extend(c.tolist())
                n_cats = len(dim.categories)
                dist = p0 = p1 = loc = scale = p2 = 0
                low, high = 0.0, float(n_cats)
            else:
                dist, p0, p1, loc, scale, p2 = (tuple(_prior_params(dim)) + (0.0,))[:6]
                if dist == _lib.DIST_NONE and self.unsamplable is None:
                    self.unsamplable = (dim.name, dim._prior_name)
                low, high = dim.interval()
                low, high = float(low), float(high)
                cdf_off = n_cats = 0
            self.dim_bounds.append((low, high))
            self.dim_dists.append(None if kind == 'categorical' else dist)
            self.dim_generic_rvs.append(kind != 'categorical' and _draws_by_inverse_cdf(dim.prior))
            op = _OPS[requirement][kind]
            if op == _lib.OP_ONEHOT:
                width, t_elem, t_dt = (1 if n_cats <= 2 else n_cats), 8, 'f8'
            elif op in (_lib.OP_QUANTIZE,):
                width, t_elem, t_dt = 1, 8, 'i8'
            elif op == _lib.OP_TO_REAL:
                width, t_elem, t_dt = 1, 8, 'f8'
            else:
                width = 1
                t_elem = 4 if kind == 'categorical' else 8
                t_dt = {'real': 'f8', 'integer': 'i8', 'categorical': 'i4'}[kind]
            elem = 4 if kind == 'categorical' else 8
            dt = {'real': 'f8', 'integer': 'i8', 'categorical': 'i4'}[kind]
            for e in range(k):
                sd = SlotDesc()
                s = len(slots)
                sd.kind, sd.dist = _KIND[kind], dist
                sd.dim_first = first
                sd.flags = _lib.SLOT_DIM_LAST if e == k - 1 else 0
                if kind == 'categorical' and len(set(dim._probs)) == 1:
                    sd.flags |= _lib.SLOT_CAT_UNIFORM
                sd.n_cats, sd.cdf_off = n_cats, cdf_off
                sd.col_units, sd.t_units = col_units, t_units
                sd.t_op = op
                sd.t_flags = _lib.TFLAG_FLOOR if quantize == 'floor' else 0
                sd.t_width = width
                sd.p0, sd.p1, sd.loc, sd.scale = p0, p1, loc, scale
                sd.low, sd.high = low, high
                for j, v in enumerate(_hash_keys(s)):
                    sd.hk[j] = v
                sd.hk[3] = int(np.float64(p2).view(np.uint64))      # third prior parameter (beta: lbeta(a, b))
                slots.append(sd)
                self.col_meta.append((col_units, dt))
                for j in range(width):
                    self.tcol_meta.append((t_units + j * t_elem, t_dt, s))
                col_units += elem
                t_units += t_elem * width
        self.n_slots = len(slots)
        self.dim_widths = [b - a for a, b in zip(self.dim_first, self.dim_first[1:] + [len(slots)])]
        self._col_runs = []                                 # (first column's units, dtype, number of columns)
        for units, dt in self.col_meta:
            if self._col_runs and self._col_runs[-1][1] == dt:
                self._col_runs[-1][2] += 1
            else:
                self._col_runs.append([units, dt, 1])
        self.n_tcols = len(self.tcol_meta)
        self.row_bytes = col_units
        self.trow_bytes = t_units
        arr = (SlotDesc * self.n_slots)(*slots)
        cdf_arr = (ctypes.c_double * max(1, len(cdf)))(*cdf)
        handle = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.ob2_space_create(arr, self.n_slots, cdf_arr, len(cdf), ctypes.byref(handle)))
        self.handle = handle
        self.philox_rounds = int(philox_rounds)
        if self.philox_rounds != 10:                    # 7: Philox4x32-7 (fewer instructions, other draws)
            _lib.check(self.lib.ob2_space_set_philox_rounds(handle, self.philox_rounds))
        self._host_cols = self._host_failed = None      # sample_host's reusable host block
        assert self.lib.ob2_space_row_bytes(handle) == self.row_bytes
        assert self.lib.ob2_space_trow_bytes(handle) == self.trow_bytes
        # the narrowed host hand-off layout (ob2_compact_wire): the library decides it, we only read it
        units = (ctypes.c_int32 * self.n_slots)()
        widths = (ctypes.c_int32 * self.n_slots)()
        _lib.check(self.lib.ob2_space_wire_layout(handle, units, widths))
        self.wire_row_bytes = int(self.lib.ob2_space_wire_row_bytes(handle))
        self.wire_meta = []
        for s, (_u, dt) in enumerate(self.col_meta):
            w = int(widths[s])
            wdt = 'f8' if dt == 'f8' else ('i%d' % w) if dt == 'i8' else ('u%d' % w if w < 4 else 'i4')
            self.wire_meta.append((int(units[s]), wdt))

    def __del__(self):
        handle = getattr(self, 'handle', None)
        if handle:
            try:
                self.lib.ob2_space_destroy(handle)
            except Exception:      # interpreter shutdown
                pass
            self.handle = None

    # ------------------------------------------------------------ buffers
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def empty_columns(self, n, transformed=False, wire=False):
        """Uninitialised column block for n candidates (``wire``: the narrowed
        host hand-off layout of ``ob2_compact_wire``)."""
        from .columns import ColumnBatch
        assert not (transformed and wire)
        n_stride = round_up(max(int(n), 1))
        row = self.wire_row_bytes if wire else self.trow_bytes if transformed else self.row_bytes
        buf = torch.empty(row * n_stride, dtype=torch.uint8, device=self.device)
        return ColumnBatch(self, int(n), n_stride, buf, transformed, wire=wire)

    # ------------------------------------------------------------ kernels
    def sample(self, n, seed, base_index=0, with_mask=True, with_fingerprint=False, out=None, index=None,
               transformed_out=None):
        """Philox-mode Space.sample: returns a ColumnBatch (device).  ``index``
        (int64 device tensor [n]) replaces the consecutive counters.  With
        ``transformed_out`` (a transformed ColumnBatch of the same stride) the
        fused kernel also writes transform(sample) into it."""
        self._require_samplable()
        batch = out if out is not None else self.empty_columns(n)
        assert not batch.transformed and batch.n == n
        ns = batch.n_stride
        if transformed_out is not None:
            tb = transformed_out
            assert tb.transformed and tb.n_stride == ns and tb.compiled is self
            if with_mask and batch.mask is None:
                batch.mask = torch.empty(ns, dtype=torch.uint8, device=self.device)
            if with_fingerprint and batch.fp is None:
                batch.fp = torch.empty((ns, 2), dtype=torch.int64, device=self.device)
            if batch.n_failed is None:
                batch.n_failed = torch.zeros(self.n_slots, dtype=torch.int32, device=self.device)
            else:
                batch.n_failed.zero_()
            _lib.check(self.lib.ob2_sample_transform(
                self.handle, n, ns, int(seed) & _M64, int(base_index) & _M64,
                index.data_ptr() if index is not None else None, batch.buf.data_ptr(), tb.buf.data_ptr(),
                batch.mask.data_ptr() if with_mask else None,
                batch.fp.data_ptr() if with_fingerprint else None,
                batch.n_failed.data_ptr(), self._stream()))
            tb.n = n
            return batch
        if with_mask and batch.mask is None:
            batch.mask = torch.empty(ns, dtype=torch.uint8, device=self.device)
        if with_fingerprint and batch.fp is None:
            batch.fp = torch.empty((ns, 2), dtype=torch.int64, device=self.device)
        if batch.n_failed is None:
            batch.n_failed = torch.zeros(self.n_slots, dtype=torch.int32, device=self.device)
        else:
            batch.n_failed.zero_()
        _lib.check(self.lib.ob2_sample(
            self.handle, n, ns, int(seed) & _M64, int(base_index) & _M64,
            index.data_ptr() if index is not None else None, batch.buf.data_ptr(),
            batch.mask.data_ptr() if with_mask else None,
            batch.fp.data_ptr() if with_fingerprint else None,
            batch.n_failed.data_ptr(), self._stream()))
        return batch

    def suggest(self, n, seed, base_index=0, index=None, out=None, transformed_out=None, back_out=None):
        """Random.suggest behind PrimaryAlgo as one kernel (``ob2_suggest``):
        sample -> transform -> reverse -> fingerprint.  Returns ``(batch, tbatch,
        back)``; ``back.fp`` holds the fingerprints of the handed-out rows and
        ``batch.mask`` the fused ``point in space``.  Bit-identical to
        ``sample(transformed_out=...)`` followed by ``reverse(with_fingerprint=True)``.
        ``back_out`` may be the same batch as ``out``: only the handed-out rows
        are then kept (one store per value instead of two)."""
        self._require_samplable()
        batch = out if out is not None else self.empty_columns(n)
        tb = transformed_out if transformed_out is not None else self.empty_columns(n, transformed=True)
        back = back_out if back_out is not None else self.empty_columns(n)
        ns = batch.n_stride
        assert not batch.transformed and tb.transformed and not back.transformed
        assert batch.n == n and tb.n_stride == ns and back.n_stride == ns
        if batch.mask is None:
            batch.mask = torch.empty(ns, dtype=torch.uint8, device=self.device)
        if back.fp is None:
            back.fp = torch.empty((ns, 2), dtype=torch.int64, device=self.device)
        if batch.n_failed is None:
            batch.n_failed = torch.zeros(self.n_slots, dtype=torch.int32, device=self.device)
        else:
            batch.n_failed.zero_()
        _lib.check(self.lib.ob2_suggest(
            self.handle, n, ns, int(seed) & _M64, int(base_index) & _M64,
            index.data_ptr() if index is not None else None, batch.buf.data_ptr(), tb.buf.data_ptr(),
            back.buf.data_ptr(), batch.mask.data_ptr(), back.fp.data_ptr(), batch.n_failed.data_ptr(),
            self._stream()))
        tb.n = n
        back.n = n
        return batch, tb, back

    def sample_host(self, n, seed, base_index=0, history=None):
        """Small draws (n <= ``_lib.HOST_SAMPLE_MAX``) delivered as host columns
        by ONE library call (``ob2_sample_host``: kernel, copy and sync inside).
        Returns (list of numpy columns, n_failed numpy array); with ``history``
        (a TrialHistory) the rows are also deduplicated against it
        (``ob2_sample_dedup_host``) and a bool keep array is returned third."""
        self._require_samplable()
        n = int(n)
        ns = round_up(max(n, 1))
        if self._host_cols is None:
            self._host_cols = np.empty(self.row_bytes * _lib.HOST_SAMPLE_MAX, dtype=np.uint8)
            self._host_failed = np.empty(self.n_slots, dtype=np.uint32)
        host = self._host_cols
        keep = None
        if history is not None:         # draw + fingerprint + dedup against the history, still one call
            keep = np.empty(max(n, 1), dtype=np.uint8)
            rc = self.lib.ob2_sample_dedup_host(self.handle, history.handle, n, int(seed) & _M64,
                                                int(base_index) & _M64, host.ctypes.data, keep.ctypes.data,
                                                self._host_failed.ctypes.data, self._stream())
            if rc == -4 and history.auto_grow:      # set full (refused before anything was touched)
                history.grow()
                return self.sample_host(n, seed, base_index, history)
            _lib.check(rc)
            keep = keep[:n].astype(bool)
        else:
            _lib.check(self.lib.ob2_sample_host(self.handle, n, int(seed) & _M64, int(base_index) & _M64,
                                                host.ctypes.data, self._host_failed.ctypes.data, self._stream()))
        blk = host[:self.row_bytes * ns].copy()             # the caller's own copy; the columns are views of it
        cols = []
        for units, dt, count in self._col_runs:             # adjacent columns of one dtype: one 2-D view per run
            e = _ELEM[dt]
            cols.extend(blk[units * ns:(units + e * count) * ns].view(dt).reshape(count, ns)[:, :n])
        if history is not None:
            return cols, self._host_failed, keep
        return cols, self._host_failed

    def sample_from_base_host(self, base, n, normal_variates=True):
        """``base``: float64 numpy [n_slots][n] of the caller's random numbers (n <= ``_lib.HOST_SAMPLE_MAX``)
        -> (list of numpy columns, bool ok[n]) by ONE library call (``ob2_sample_from_base_host``)."""
        self._require_samplable()
        n = int(n)
        ns = round_up(max(n, 1))
        block = np.full((self.n_slots, ns), 0.5)
        block[:, :n] = base
        blk = np.empty(self.row_bytes * ns, dtype=np.uint8)
        ok = np.empty(max(n, 1), dtype=np.uint8)
        _lib.check(self.lib.ob2_sample_from_base_host(
            self.handle, n, block.ctypes.data, _lib.BASE_NORMAL_VARIATE if normal_variates else 0,
            blk.ctypes.data, ok.ctypes.data, self._stream()))
        cols = []
        for units, dt, count in self._col_runs:
            e = _ELEM[dt]
            cols.extend(blk[units * ns:(units + e * count) * ns].view(dt).reshape(count, ns)[:, :n])
        return cols, ok[:n].astype(bool)

    def contains_host(self, cols, n, int_as_f64=True):
        """``point in space`` for n <= ``_lib.HOST_SAMPLE_MAX`` host rows given as
        columns (Integer columns as float64 when ``int_as_f64``), by ONE library
        call (``ob2_contains_host``).  Returns a bool numpy array [n]."""
        ns = round_up(max(int(n), 1))
        host = np.zeros(self.row_bytes * ns, dtype=np.uint8)
        for (units, dt), col in zip(self.col_meta, cols):
            want = np.float64 if (int_as_f64 and dt == 'i8') else dt
            arr = np.ascontiguousarray(col, dtype=want)
            host[units * ns:units * ns + arr.nbytes] = arr.view(np.uint8)
        mask = np.empty(max(int(n), 1), dtype=np.uint8)
        _lib.check(self.lib.ob2_contains_host(self.handle, int(n), host.ctypes.data, 1 if int_as_f64 else 0,
                                              mask.ctypes.data, self._stream()))
        return mask[:n].astype(bool)

    def _require_samplable(self):
        if self.unsamplable is not None:
            raise NotImplementedError(
                "Dimension '{}': prior '{}' has no device inverse CDF yet (supported: uniform, reciprocal / "
                "loguniform, norm and the eighty-three further scipy priors INTEGRATION.md lists). "
                "There is no CPU fallback.".format(*self.unsamplable))

    def sample_from_uniform(self, U, n=None, with_fingerprint=False, normal_variates=False):
        """Injected random stream: U is a float64 tensor [n_slots][n_stride] (device) of uniforms -- the
        parity mode -- or, with ``normal_variates``, of what numpy's RandomState hands scipy per draw
        (standard normal variates for ``norm`` slots; ``ob2_sample_from_base``, see :mod:`.mt_stream`)."""
        self._require_samplable()
        assert U.dtype == torch.float64 and U.is_cuda and U.dim() == 2 and U.shape[0] == self.n_slots
        ns = U.shape[1]
        if ns % 32:
            raise ValueError("U must be padded to a multiple of 32 columns")
        n = ns if n is None else int(n)
        batch = self.empty_columns(ns)
        batch.n = n
        batch.mask = torch.empty(ns, dtype=torch.uint8, device=self.device)
        if with_fingerprint:
            batch.fp = torch.empty((ns, 2), dtype=torch.int64, device=self.device)
        _lib.check(self.lib.ob2_sample_from_base(
            self.handle, n, ns, U.data_ptr(), _lib.BASE_NORMAL_VARIATE if normal_variates else 0,
            batch.buf.data_ptr(), batch.mask.data_ptr(),
            batch.fp.data_ptr() if with_fingerprint else None, self._stream()))
        return batch

    def contains(self, batch, int_as_f64=False):
        """Space.__contains__ over a ColumnBatch -> uint8 tensor [n]."""
        assert not batch.transformed
        mask = torch.empty(batch.n_stride, dtype=torch.uint8, device=self.device)
        _lib.check(self.lib.ob2_contains(self.handle, batch.n, batch.n_stride, batch.buf.data_ptr(),
                                         1 if int_as_f64 else 0, mask.data_ptr(), self._stream()))
        return mask[:batch.n]

    def transform(self, batch, out=None):
        assert not batch.transformed
        tb = out if out is not None else self.empty_columns(batch.n, transformed=True)
        assert tb.transformed and tb.n_stride == batch.n_stride
        _lib.check(self.lib.ob2_transform(self.handle, batch.n, batch.n_stride, batch.buf.data_ptr(),
                                          tb.buf.data_ptr(), self._stream()))
        return tb

    def reverse(self, tbatch, out=None, with_fingerprint=False):
        """Transformed batch -> original batch.  ``with_fingerprint`` also fills
        ``out.fp`` with the fingerprints of the rows written (fused into the
        same bandwidth-bound kernel)."""
        assert tbatch.transformed
        b = out if out is not None else self.empty_columns(tbatch.n)
        assert not b.transformed and b.n_stride == tbatch.n_stride
        b.n = tbatch.n
        if with_fingerprint:
            if b.fp is None:
                b.fp = torch.empty((b.n_stride, 2), dtype=torch.int64, device=self.device)
            _lib.check(self.lib.ob2_reverse_fingerprint(self.handle, tbatch.n, tbatch.n_stride, tbatch.buf.data_ptr(),
                                                        b.buf.data_ptr(), b.fp.data_ptr(), self._stream()))
        else:
            _lib.check(self.lib.ob2_reverse(self.handle, tbatch.n, tbatch.n_stride, tbatch.buf.data_ptr(),
                                            b.buf.data_ptr(), self._stream()))
        return b

    def fingerprint(self, batch):
        assert not batch.transformed
        if batch.fp is None:
            batch.fp = torch.empty((batch.n_stride, 2), dtype=torch.int64, device=self.device)
        _lib.check(self.lib.ob2_fingerprint(self.handle, batch.n, batch.n_stride, batch.buf.data_ptr(),
                                            batch.fp.data_ptr(), self._stream()))
        return batch.fp


_CACHE_ATTR = '_ob2_compiled'


def compile_space(space, requirement=None, quantize='round', device=None, philox_rounds=10):
    """Compile (and memoise on the Space object) for the given requirement."""
    key = (requirement, quantize, str(device) if device is not None else torch.cuda.current_device(),
           tuple(id(d) for d in space.values()), philox_rounds)
    cache = space.__dict__.setdefault(_CACHE_ATTR, {}) if hasattr(space, '__dict__') else {}
    hit = cache.get(key)
    if hit is None:
        hit = CompiledSpace(space, requirement, quantize, device, philox_rounds)
        cache[key] = hit
    return hit
