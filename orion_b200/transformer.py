"""TransformedSpace and the transformer chain, executed on the device.

The reference snapshot only has TODO markers where this module belongs
(``src/orion/core/worker/primary_algo.py:13-14,41-42,54-55``); the names and
semantics follow upstream Orion's ``orion.core.worker.transformer`` as restated
in SURVEY.md Appendix B (parity unpinned by /root/reference).

The classes below *describe* a chain (types, target shapes, intervals and the
device op it lowers to); the arithmetic itself runs in ``ob2_transform`` /
``ob2_reverse`` for whole batches.  Per-point calls go through the same kernels
with a batch of one -- there is no host implementation to fall back on.
"""
from collections import OrderedDict

import numpy

from . import _lib
from .columns import ColumnBatch
from .descriptor import _OPS, compile_space
from .space import Space, _raise_if_improbable, philox_key


class Transformer(object):
    """Describes one step of a chain: ``domain_type`` -> ``target_type``."""
    domain_type = None
    target_type = None

    def infer_target_shape(self, shape):
        return shape

    def interval(self, low, high):
        return (low, high)

    def __repr__(self):
        return "{}()".format(type(self).__name__)


class Identity(Transformer):
    """No-op; keeps whatever type it is given."""

    def __init__(self, domain_type=None):
        self.domain_type = self.target_type = domain_type


class Quantize(Transformer):
    """real -> integer by ``numpy.round`` (half to even); ``reverse`` casts
    back to float."""
    domain_type = 'real'
    target_type = 'integer'


class Enumerate(Transformer):
    """categorical -> integer: a category becomes its position in
    ``categories``; interval ``(0, len(categories))``."""
    domain_type = 'categorical'
    target_type = 'integer'

    def __init__(self, categories):
        self.categories = tuple(categories)

    def interval(self, low=None, high=None):
        return (0, len(self.categories))

    def __repr__(self):
        return "Enumerate({} categories)".format(len(self.categories))


class OneHotEncode(Transformer):
    """integer -> real: index i of ``num_cats`` becomes a trailing axis of
    length ``num_cats`` with 1.0 at i (a single float when num_cats <= 2);
    ``reverse`` is ``argmax`` (``> 0.5`` for two categories)."""
    domain_type = 'integer'
    target_type = 'real'

    def __init__(self, num_cats):
        self.num_cats = int(num_cats)

    def infer_target_shape(self, shape):
        return tuple(shape) + (self.num_cats,) if self.num_cats > 2 else tuple(shape)

    def interval(self, low=None, high=None):
        if self.num_cats <= 2:
            return (0.0, 1.0)
        return ((0.0,) * self.num_cats, (1.0,) * self.num_cats)

    def __repr__(self):
        return "OneHotEncode({})".format(self.num_cats)


class Reverse(Transformer):
    """Swap ``transform`` and ``reverse`` of the wrapped transformer."""

    def __init__(self, transformer):
        self.transformer = transformer
        self.domain_type = transformer.target_type
        self.target_type = transformer.domain_type

    def __repr__(self):
        return "Reverse({!r})".format(self.transformer)


class Compose(Transformer):
    """``transform = t_n o ... o t_1``; types chain end to end."""

    def __init__(self, transformers, base_domain_type=None):
        self.transformers = list(transformers)
        self.domain_type = self.transformers[0].domain_type if self.transformers else base_domain_type
        self.target_type = self.transformers[-1].target_type if self.transformers else base_domain_type

    def infer_target_shape(self, shape):
        for t in self.transformers:
            shape = t.infer_target_shape(shape)
        return shape

    def interval(self, low, high):
        out = (low, high)
        for t in self.transformers:
            out = t.interval(*out)
        return out

    def __repr__(self):
        return "Compose({!r})".format(self.transformers)


def _device_op(transformer, kind):
    """Lower a chain to the kernel's per-slot op code."""
    t = transformer
    if isinstance(t, Compose):
        steps = t.transformers
        if not steps:
            return _lib.OP_IDENTITY
        if len(steps) == 1:
            return _device_op(steps[0], kind)
        if len(steps) == 2 and isinstance(steps[0], Enumerate) and isinstance(steps[1], OneHotEncode):
            return _lib.OP_ONEHOT
        raise NotImplementedError("no device lowering for {!r}".format(t))
    if isinstance(t, Identity):
        return _lib.OP_IDENTITY
    if isinstance(t, Quantize):
        return _lib.OP_QUANTIZE
    if isinstance(t, Reverse) and isinstance(t.transformer, Quantize):
        return _lib.OP_TO_REAL
    if isinstance(t, Enumerate):
        return _lib.OP_ENUMERATE
    raise NotImplementedError("no device lowering for {!r}".format(t))


_REQUIRED = {
    'real': {'real': lambda d: Identity('real'),
             'integer': lambda d: Reverse(Quantize()),
             'categorical': lambda d: Compose([Enumerate(d.categories), OneHotEncode(len(d.categories))])},
    'integer': {'real': lambda d: Quantize(),
                'integer': lambda d: Identity('integer'),
                'categorical': lambda d: Enumerate(d.categories)},
    'numerical': {'real': lambda d: Identity('real'),
                  'integer': lambda d: Identity('integer'),
                  'categorical': lambda d: Compose([Enumerate(d.categories), OneHotEncode(len(d.categories))])},
    None: {'real': lambda d: Identity('real'),
           'integer': lambda d: Identity('integer'),
           'categorical': lambda d: Identity('categorical')},
}


class TransformedDimension(object):
    """An original dimension seen through a transformer."""

    def __init__(self, transformer, original_dimension, requirement=None, quantize='round'):
        self.transformer = transformer
        self.original_dimension = original_dimension
        self._requirement = requirement
        self._quantize = quantize
        self._solo = None

    def _solo_space(self):
        """This dimension alone as a one-dimension TransformedSpace (the device
        path works on spaces); built on first use."""
        if self._solo is None:
            solo = Space()
            OrderedDict.__setitem__(solo, self.original_dimension.name or '_', self.original_dimension)
            self._solo = TransformedSpace(solo, self._requirement, self._quantize)
        return self._solo

    @property
    def name(self):
        return self.original_dimension.name

    @property
    def type(self):
        return self.transformer.target_type or self.original_dimension.type

    @property
    def shape(self):
        return tuple(self.transformer.infer_target_shape(self.original_dimension.shape))

    def interval(self, alpha=1.0):
        if self.original_dimension.type == 'categorical':
            return self.transformer.interval(None, None)
        return self.transformer.interval(*self.original_dimension.interval(alpha))

    def transform(self, point):
        return self._solo_space().transform((point,))[0]

    def reverse(self, transformed_point):
        return self._solo_space().reverse((transformed_point,))[0]

    def sample(self, n_samples=1, seed=None):
        return [self.transform(p) for p in self.original_dimension.sample(n_samples, seed)]

    def __contains__(self, point):
        return self.reverse(point) in self.original_dimension

    def __repr__(self):
        return "{!r} <- {!r}".format(self.transformer, self.original_dimension)


class TransformedSpace(Space):
    """A Space of :class:`TransformedDimension` with whole-point and
    whole-batch ``transform`` / ``reverse``."""

    def __init__(self, original_space, requirement, quantize='round'):
        super(TransformedSpace, self).__init__()
        self.original = original_space
        self.requirement = requirement
        self.quantize = quantize
        for dim in original_space.values():
            kind = 'real' if dim.type == 'dimension' else dim.type
            t = _REQUIRED[requirement][kind](dim)
            if _device_op(t, kind) != _OPS[requirement][kind]:
                raise NotImplementedError("no device lowering for {!r}".format(t))
            # (Space.__setitem__ only admits Dimension instances)
            OrderedDict.__setitem__(self, dim.name, TransformedDimension(t, dim, requirement, quantize))

    def compiled(self, requirement=None, quantize=None, philox_rounds=10):
        return compile_space(self.original, self.requirement, self.quantize, philox_rounds=philox_rounds)

    # -- batches -------------------------------------------------------------
    def transform_columns(self, batch, out=None):
        """Original ColumnBatch -> transformed ColumnBatch (device)."""
        return self.compiled().transform(batch, out=out)

    def reverse_columns(self, tbatch, out=None):
        """Transformed ColumnBatch -> original ColumnBatch (device)."""
        return self.compiled().reverse(tbatch, out=out)

    def sample_columns(self, n_samples=1, seed=None, base_index=0, check=True, **_kw):
        """``transform(original.sample(...))`` without leaving the device;
        returns (original batch, transformed batch)."""
        compiled = self.compiled()
        tbatch = compiled.empty_columns(n_samples, transformed=True)
        batch = compiled.sample(n_samples, philox_key(seed), base_index=base_index, with_mask=True,
                                transformed_out=tbatch)          # one fused kernel
        if check:
            _raise_if_improbable(batch, list(self.original.values()))
        return batch, tbatch

    def suggest_columns(self, n_samples=1, seed=None, base_index=0, check=True):
        """What ``Random`` behind ``PrimaryAlgo`` produces, from one kernel
        (``ob2_suggest``): returns (original batch, transformed batch, handed-out
        batch = reverse(transformed)); the last carries the row fingerprints
        (``.fp``) the history dedups on, the first the fused in-space mask."""
        compiled = self.compiled()
        batch, tbatch, back = compiled.suggest(n_samples, philox_key(seed), base_index=base_index)
        if check:
            _raise_if_improbable(batch, list(self.original.values()))
        return batch, tbatch, back

    def contains_columns(self, tbatch):
        """``reverse(p) in original`` for every transformed row."""
        compiled = self.compiled()
        return compiled.contains(compiled.reverse(tbatch))

    # -- points (compatibility path) ----------------------------------------
    def _t_widths(self):
        """Transformed columns per original dimension."""
        c = self.compiled()
        per_slot = {}
        for _u, _dt, s in c.tcol_meta:
            per_slot[s] = per_slot.get(s, 0) + 1
        widths = []
        for j, first in enumerate(c.dim_first):
            last = c.dim_first[j + 1] if j + 1 < len(c.dim_first) else c.n_slots
            widths.append([per_slot[s] for s in range(first, last)])
        return widths

    def transform_points(self, points):
        """List of original tuples -> list of transformed tuples."""
        cols, row_ok = self.original.columns_from_rows(points)
        if not row_ok.all():
            raise ValueError("point does not match the shape of the original space")
        compiled = self.compiled()
        # Integer columns were produced as float64 by columns_from_rows; the
        # transform kernels read the typed layout
        typed = [numpy.asarray(c, dtype={'f8': numpy.float64, 'i8': numpy.int64, 'i4': numpy.int32}[dt])
                 for c, (_u, dt) in zip(cols, compiled.col_meta)]
        batch = ColumnBatch.from_numpy(compiled, typed)
        if not bool(compiled.contains(batch).all().item()):
            raise AssertionError("point outside the original space")   # OneHotEncode/Enumerate assert their domain
        tcols = compiled.transform(batch).to_numpy()
        return self._rows_from_tcols(tcols, len(points))

    def _rows_from_tcols(self, tcols, n):
        per_dim = []
        t = 0
        for tdim, slot_widths in zip(self.values(), self._t_widths()):
            k = sum(slot_widths)
            block = tcols[t:t + k]
            t += k
            shape = tdim.shape
            if shape == ():
                per_dim.append([block[0][i] for i in range(n)])
            else:
                per_dim.append([numpy.stack([c[i] for c in block]).reshape(shape) for i in range(n)])
        return list(zip(*per_dim))

    def reverse_points(self, tpoints):
        """List of transformed tuples -> list of original tuples."""
        compiled = self.compiled()
        n = len(tpoints)
        tcols = []
        for j, tdim in enumerate(self.values()):
            for e in range(int(numpy.prod(tdim.shape, dtype=numpy.int64)) if tdim.shape else 1):
                tcols.append([])
        for point in tpoints:
            t = 0
            for j, tdim in enumerate(self.values()):
                arr = numpy.asarray(point[j])
                if arr.shape != tdim.shape:
                    raise ValueError("transformed point does not match the shape of dimension '{}'".format(tdim.name))
                for v in arr.reshape(-1):
                    tcols[t].append(v)
                    t += 1
        typed = [numpy.asarray(c, dtype={'f8': numpy.float64, 'i8': numpy.int64, 'i4': numpy.int32}[dt])
                 for c, (_u, dt, _s) in zip(tcols, compiled.tcol_meta)]
        tbatch = ColumnBatch.from_numpy(compiled, typed, transformed=True)
        cols = compiled.reverse(tbatch).to_numpy()
        return self.original.rows_from_columns(cols, n)

    def transform(self, point):
        return self.transform_points([tuple(point)])[0]

    def reverse(self, transformed_point):
        return self.reverse_points([tuple(transformed_point)])[0]

    def sample(self, n_samples=1, seed=None):
        """``[transform(p) for p in original.sample(n, seed)]`` -- computed as
        one device batch."""
        cols = self.original.reference_stream_columns(n_samples, seed)
        if cols is not None:                        # the reference's own stream, then the transform kernel
            batch = ColumnBatch.from_numpy(self.compiled(), cols)
            return self._rows_from_tcols(self.compiled().transform(batch).to_numpy(), n_samples)
        _, tbatch = self.sample_columns(n_samples, seed)
        return self._rows_from_tcols(tbatch.to_numpy(), n_samples)

    def contains_points(self, tpoints):
        try:
            originals = self.reverse_points(tpoints)
        except ValueError:
            return [False] * len(tpoints)
        return self.original.contains_points(originals)

    def __contains__(self, value):
        if isinstance(value, str):
            return dict.__contains__(self, value)
        try:
            len(value)
        except TypeError as exc:
            raise TypeError("Can check only for dimension names or "
                            "for tuples with parameter values.") from exc
        if not self:
            return False
        return bool(self.contains_points([tuple(value)])[0])

    def interval(self, alpha=1.0):
        return [tdim.interval(alpha) for tdim in self.values()]


def build_required_space(requirement, original_space, quantize='round'):
    """TransformedSpace meeting an algorithm's requirement
    ('real' | 'integer' | 'numerical' | None); SURVEY.md Appendix B.2."""
    if requirement not in _REQUIRED:
        raise TypeError("Unsupported space requirement: {}".format(requirement))
    return TransformedSpace(original_space, requirement, quantize)
