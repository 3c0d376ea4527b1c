"""Search-space objects with the reference's interface, backed by the device.

Host-side mirror of ``src/orion/algo/space.py`` (reference): the same class
names (`Dimension`, `Real`, `Integer`, `Categorical`, `Space`), constructor
arguments, error behaviour and return types, so code written against
``orion.algo.space`` runs unchanged.  What differs is where the work happens:

* ``sample`` computes on the GPU, not through ``scipy.rvs`` one scalar at a
  time.  The list-returning calls (``Space.sample``, ``Dimension.sample``)
  feed the kernel the reference's own MT19937 stream (:mod:`.mt_stream`), so
  a seed (None / int / RandomState) returns the reference's values and leaves
  the RandomState where the reference leaves it -- for the priors whose scipy
  ``_rvs`` is a function of ``random_sample`` / ``standard_normal`` numbers
  (uniform, reciprocal / loguniform, norm, Categorical).  The tensor calls
  (``sample_columns`` and everything bulk) and the remaining priors draw from
  Philox4x32 keyed by the seed (inverse CDF, fused bounds check, <= 4 attempts):
  reproducible under the same seeding discipline, other values.
* ``in`` checks run the ``ob2_contains`` kernel with bounds resolved once.
* ``Space.sample_columns`` / ``Space.contains_columns`` are the added
  tensor-returning fast path; ``Space.sample`` converts the same device result
  to the reference's list of tuples.

scipy is used on the host only for what is static per dimension (argument
parsing, ``interval``), exactly the calls the reference makes.
"""
from collections import OrderedDict

import numpy
from scipy._lib._util import check_random_state
from scipy.stats import distributions

from . import mt_stream
from ._lib import HOST_SAMPLE_MAX
from .columns import ColumnBatch
from .descriptor import compile_space

_BANNED = {
    'random_state': "random_state/seed cannot be set in a parameter's definition! Set seed globally!",
    'seed': "random_state/seed cannot be set in a parameter's definition! Set seed globally!",
    'discrete': "Do not use kwarg 'discrete' on `Dimension`, use pure `_Discrete` class instead!",
    'size': "Use 'shape' keyword only instead of 'size'.",
}


def _stream_owner(seed):
    """The RandomState a seed argument stands for (an int seeds fresh ones: any RandomState will do
    to say so), or whatever else ``check_random_state`` makes of it."""
    if isinstance(seed, (int, numpy.integer)):
        return _ANY_RANDOM_STATE
    return check_random_state(seed)


_ANY_RANDOM_STATE = numpy.random.RandomState(0)


def _solo_compiled(dim):
    """``dim`` compiled as a one-dimension space, cached on the object (also a reference Dimension)."""
    compiled = dim.__dict__.get('_compiled_self')
    if compiled is None:
        solo = _OneDim()
        solo[dim.name or '_'] = dim
        compiled = dim.__dict__['_compiled_self'] = compile_space(solo)
    return compiled


def philox_key(seed):
    """64-bit Philox key from the reference's ``seed`` argument
    (None -> numpy's global RNG, int -> a fresh RandomState(int),
    RandomState -> drawn from it, advancing it; space.py:118-127, 482)."""
    rng = check_random_state(seed)
    if hasattr(rng, 'randint'):
        w = rng.randint(0, 2 ** 32, size=2, dtype=numpy.uint64)
    else:                       # numpy.random.Generator
        w = rng.integers(0, 2 ** 32, size=2, dtype=numpy.uint64)
    return int(w[0]) | (int(w[1]) << 32)


class _OneDim(OrderedDict):
    """A throw-away single-dimension space used to compile one Dimension."""


# ---- value <-> column helpers.  Module-level and duck-typed (they read only ``type``, ``shape``
# and ``categories``) so that they also serve the REFERENCE's own dimension objects when a
# reference ``Space`` is driven through the kernels (compat / plugin).
def _index_of(dim, value):
    """Position of ``value`` in ``dim.categories`` by ``==`` (so '2' is not 2), -1 when absent."""
    try:
        return tuple(dim.categories).index(value)
    except ValueError:
        return -1


_MINUS_ONES = __import__('itertools').repeat(-1)


def _layout_of(space):
    """``([dim.shape ...], [dim.type ...])`` of a space, remembered on it (scipy parses the prior's arguments
    for every ``shape`` of a reference Dimension; the key notices a dimension registered since)."""
    key = tuple(map(id, space.values()))
    hit = space.__dict__.get('_ob2_layout')
    if hit is None or hit[0] != key:
        dims = list(space.values())
        hit = space.__dict__['_ob2_layout'] = (key, [tuple(d.shape) for d in dims], [d.type for d in dims])
    return hit[1], hit[2]


def _category_lookup(dim):
    """category -> first index with an equal category (what ``tuple.index`` finds: 2 == 2.0, '2' != 2), cached on
    the dimension; None when some category is unhashable."""
    lookup = dim.__dict__.get('_ob2_lookup', False)
    if lookup is False:
        lookup = {}
        try:
            for i, c in enumerate(dim.categories):
                lookup.setdefault(c, i)
        except TypeError:
            lookup = None
        dim.__dict__['_ob2_lookup'] = lookup
    return lookup


def _column_values(dim, point):
    """Flatten one value of ``dim`` into per-slot scalars, or None when its shape does not fit
    (``point_.shape != self.shape`` -> False, space.py:162,420).  A string offered to a numeric
    dimension raises TypeError, as the reference's ``point_ < high`` does (space.py:164)."""
    shape = tuple(dim.shape)
    if dim.type == 'categorical':
        arr = numpy.asarray(point, dtype=object)
        if arr.shape != shape:
            return None
        return numpy.asarray([_index_of(dim, v) for v in arr.reshape(-1)], dtype=numpy.int32)
    probe = numpy.asarray(point)
    if probe.dtype.kind in 'US':
        raise TypeError("'<' not supported between a value of dimension '{}' given as {!r} and its bounds"
                        .format(dim.name, point))
    try:
        arr = numpy.asarray(point, dtype=numpy.float64)
    except (TypeError, ValueError):
        return None
    if arr.shape != shape:
        return None
    return arr.reshape(-1)


def _from_columns(dim, cols, i):
    """Value i of ``dim`` from its host columns (numpy scalars / category objects for shape (),
    ndarrays otherwise -- what scipy's rvs and RandomState.choice hand back)."""
    shape = tuple(dim.shape)
    if dim.type == 'categorical':
        cats = dim.categories
        if shape == ():
            return cats[cols[0][i]]
        out = numpy.empty(len(cols), dtype=object)
        for e, c in enumerate(cols):
            out[e] = cats[c[i]]
        return out.reshape(shape)
    if shape == ():
        return cols[0][i]
    return numpy.stack([c[i] for c in cols]).reshape(shape)


class Dimension(object):
    """Base class for search space dimensions (reference space.py:47-196).

    Attributes
    ----------
    name : str or None
    type : str -- lower-cased class name ('real', 'integer', 'categorical')
    prior : scipy.stats rv_generic instance
    shape : tuple
    """

    def __init__(self, name, prior, *args, **kwargs):
        self._name = None
        self.name = name
        for key in ('random_state', 'seed', 'discrete'):
            if key in kwargs:
                raise ValueError(_BANNED[key])
        if isinstance(prior, str):
            self._prior_name = prior
            self.prior = getattr(distributions, prior)
        else:
            self._prior_name = prior.name
            self.prior = prior
        if 'size' in kwargs:
            raise ValueError(_BANNED['size'])
        self._args = args
        self._kwargs = kwargs
        self._shape = self._kwargs.pop('shape', None)
        self._compiled_self = None

    # -- device plumbing ---------------------------------------------------
    def _compiled(self):
        """This dimension compiled as a one-dimension space (cached)."""
        return _solo_compiled(self)

    def _column_values(self, point):
        return _column_values(self, point)

    def _from_columns(self, cols, i):
        return _from_columns(self, cols, i)

    # -- reference interface ------------------------------------------------
    def sample(self, n_samples=1, seed=None):
        """Draw ``n_samples`` values from the prior; returns a list.

        ``seed``: None | int | numpy.random.RandomState, as in the reference
        (space.py:110-136), and with the reference's values for it when the
        prior draws from ``random_sample`` / ``standard_normal`` numbers
        (:mod:`.mt_stream`): an int seeds afresh for every draw, so the same
        value comes back ``n_samples`` times; pass a RandomState to advance.
        """
        compiled = self._compiled()
        if mt_stream.reproducible(compiled, _stream_owner(seed)):
            cols = mt_stream.sample_dimension(self, compiled, n_samples, seed)
        elif n_samples <= HOST_SAMPLE_MAX:          # one library call: kernel + copy + sync
            cols, failed = compiled.sample_host(n_samples, philox_key(seed))
            _raise_if_failed(failed, compiled, [self])
        else:
            batch = compiled.sample(n_samples, philox_key(seed), with_mask=False)
            _raise_if_improbable(batch, [self])
            cols = batch.to_numpy()
        return [self._from_columns(cols, i) for i in range(n_samples)]

    def interval(self, alpha=1.0):
        """(low, high) of the prior; low inclusive, high exclusive."""
        return self.prior.interval(alpha, *self._args, **self._kwargs)

    def __contains__(self, point):
        vals = self._column_values(point)
        if vals is None:
            return False
        compiled = self._compiled()
        return bool(compiled.contains_host([numpy.asarray([v]) for v in vals], 1, int_as_f64=True)[0])

    def __repr__(self):
        return "{0}(name={1}, prior={{{2}: {3}, {4}}}, shape={5})".format(
            self.__class__.__name__, self.name, self._prior_name, self._args, self._kwargs, self.shape)

    @property
    def name(self):
        return self._name

    @name.setter
    def name(self, value):
        if value is not None and not isinstance(value, str):
            raise TypeError("Dimension's name must be either string or None. "
                            "Provided: {}, of type: {}".format(value, type(value)))
        self._name = value

    @property
    def type(self):
        return self.__class__.__name__.lower()

    @property
    def shape(self):
        # (scipy's argument parsing costs ~8 us and every conversion between points and columns asks
        # for the shape of every dimension: parsed once -- prior, arguments and shape are fixed at construction)
        shape = self.__dict__.get('_shape_parsed')
        if shape is None:
            size = self.prior._parse_args_rvs(*self._args, size=self._shape, **self._kwargs)[3]  # pylint:disable=protected-access
            shape = self.__dict__['_shape_parsed'] = tuple(int(s) for s in size)
        return shape


class Real(Dimension):
    """Real-valued dimension with optional ``low`` (inclusive) / ``high``
    (exclusive) clipping of the prior's support (reference space.py:199-276)."""

    def __init__(self, name, prior, *args, **kwargs):
        self._low = kwargs.pop('low', -numpy.inf)
        self._high = kwargs.pop('high', numpy.inf)
        if self._high <= self._low:
            raise ValueError("Lower bound {} has to be less than upper bound {}".format(self._low, self._high))
        super(Real, self).__init__(name, prior, *args, **kwargs)

    def interval(self, alpha=1.0):
        prior_low, prior_high = super(Real, self).interval(alpha)
        return (max(prior_low, self._low), min(prior_high, self._high))


class _Discrete(Dimension):
    """Floors draws and widens the interval to integers (space.py:279-319)."""

    def interval(self, alpha=1.0):
        low, high = super(_Discrete, self).interval(alpha)
        int_low = int(numpy.floor(low)) if numpy.isfinite(low) else -numpy.inf
        int_high = int(numpy.floor(high)) if numpy.isfinite(high) else numpy.inf
        if int_high < high:     # upper bound stays exclusive
            int_high += 1
        return (int_low, int_high)


class Integer(Real, _Discrete):
    """Integer dimension: ``floor`` of a draw from any prior, bounds checked on
    the floored value, membership requires integrality (space.py:322-347).
    Device dtype int64."""


class Categorical(Dimension):
    """Categorical dimension over arbitrary Python objects (space.py:350-446).

    ``categories`` is an iterable (uniform probabilities) or a dict
    category -> probability.  On the device a value is its int32 index into
    ``self.categories``.
    """

    def __init__(self, name, categories, **kwargs):
        if isinstance(categories, dict):
            self.categories = tuple(categories.keys())
            self._probs = tuple(categories.values())
        else:
            self.categories = tuple(categories)
            self._probs = tuple(numpy.tile(1. / len(self.categories), len(self.categories)))
        # rv_discrete validates that the probabilities sum to one (ValueError otherwise)
        prior = distributions.rv_discrete(values=(list(range(len(self.categories))), self._probs))
        super(Categorical, self).__init__(name, prior, **kwargs)

    def _index_of(self, value):
        return _index_of(self, value)

    def interval(self, alpha=1.0):
        raise RuntimeError("Categories have no ``interval`` (as they are not ordered).\n"
                           "Use ``self.categories`` instead.")

    def __repr__(self):
        pairs = list(zip(self.categories, self._probs))
        fmt = ['{0}: {1:.2f}'.format(c, p) for c, p in pairs]
        if len(fmt) > 5:
            fmt = fmt[:2] + ['...'] + fmt[-2:]
        return "Categorical(name={0}, prior={{{1}}}, shape={2})".format(self.name, ', '.join(fmt), self.shape)


def _raise_if_improbable(batch, dims):
    """Surface the device's rejection counters as the reference's ValueError
    (space.py:272-274).  One small device->host read."""
    _raise_if_failed(batch.n_failed.cpu().numpy(), batch.compiled, dims)


def _raise_if_failed(failed, compiled, dims):
    if not failed.any():
        return
    first_slot = int(numpy.nonzero(failed)[0][0])
    d_idx = compiled.dim_first.index(first_slot)
    dim = dims[d_idx]
    raise ValueError("Improbable bounds: (low={0}, high={1}). "
                     "Please make interval larger.".format(dim._low, dim._high))


class Space(OrderedDict):
    """Ordered mapping name -> Dimension (reference space.py:449-550)."""

    def register(self, dimension):
        self[dimension.name] = dimension

    # -- tensor fast path ---------------------------------------------------
    def compiled(self, requirement=None, quantize='round', philox_rounds=10):
        return compile_space(self, requirement, quantize, philox_rounds=philox_rounds)

    def sample_columns(self, n_samples=1, seed=None, base_index=0, with_fingerprint=False, check=True,
                       reference_stream=False):
        """Device-resident draw: returns a :class:`ColumnBatch` whose ``mask``
        is the fused ``point in space``.  ``check=False`` skips the
        synchronising read of the rejection counters.  ``reference_stream``: the
        values of the reference's ``Space.sample(n_samples, seed)`` -- its MT19937
        stream is what the kernel is fed (:mod:`.mt_stream`; the host asks numpy for
        every number, so this reproduces a reference run rather than racing it)."""
        if reference_stream:
            compiled = self.compiled()
            rng = check_random_state(seed)
            if not mt_stream.reproducible(compiled, rng):
                raise NotImplementedError("this space or seed cannot be drawn from the reference's stream "
                                          "(a prior scipy draws with another numpy generator, or a Generator seed)")
            batch = mt_stream.space_batch(list(self.values()), compiled, _solo_compiled, n_samples, rng)
            if with_fingerprint:
                compiled.fingerprint(batch)
            return batch
        batch = self.compiled().sample(n_samples, philox_key(seed), base_index=base_index,
                                       with_mask=True, with_fingerprint=with_fingerprint)
        if check:
            _raise_if_improbable(batch, list(self.values()))
        return batch

    def contains_columns(self, batch):
        """uint8 device tensor: ``point in space`` for every row of ``batch``."""
        return self.compiled().contains(batch)

    # -- reference interface ------------------------------------------------
    def sample(self, n_samples=1, seed=None):
        """List of ``n_samples`` tuples, one value per dimension in space
        order (space.py:456-484)."""
        if not self:
            return []
        cols = self.reference_stream_columns(n_samples, seed)
        if cols is not None:
            return self.rows_from_columns(cols, n_samples)
        if n_samples <= HOST_SAMPLE_MAX:            # one library call: kernel + copy + sync
            compiled = self.compiled()
            cols, failed = compiled.sample_host(n_samples, philox_key(seed))
            _raise_if_failed(failed, compiled, list(self.values()))
            return self.rows_from_columns(cols, n_samples)
        batch = self.sample_columns(n_samples, seed)
        return self.rows_from_columns(batch.to_numpy(), n_samples)

    def reference_stream_columns(self, n_samples, seed):
        """Host columns of ``[dim.sample(n, rng) for dim in self.values()]`` drawn from the reference's
        own MT19937 stream (space.py:482-484; :mod:`.mt_stream`), or None when this space or seed
        cannot be reproduced that way.  ``self`` may be a reference Space."""
        compiled = compile_space(self)
        rng = check_random_state(seed)
        if n_samples <= 0 or not mt_stream.reproducible(compiled, rng):
            return None
        return mt_stream.space_columns(list(self.values()), compiled, _solo_compiled, n_samples, rng)

    def rows_from_columns(self, host_cols, n):
        """Host columns -> the reference's list of tuples."""
        per_dim = []
        s = 0
        shapes, kinds = _layout_of(self)            # (parsed once per space; `self` may be a reference Space)
        for dim, shape, kind in zip(self.values(), shapes, kinds):
            if shape == ():                     # a whole column at once (numpy scalars, like scipy's rvs)
                col = host_cols[s][:n]
                if kind == 'categorical':
                    cats = dim.categories
                    per_dim.append([cats[c] for c in col.tolist()])
                else:
                    per_dim.append(list(col))
                s += 1
                continue
            k = int(numpy.prod(shape, dtype=numpy.int64))
            cols = host_cols[s:s + k]
            per_dim.append([_from_columns(dim, cols, i) for i in range(n)])
            s += k
        return list(zip(*per_dim))

    @staticmethod
    def _columns_of_scalar_rows(dims, points, n, layout=None):
        """The common case -- every dimension scalar, every point a full row of
        plain scalars -- column by column; None when anything is unusual (the
        general loop then decides row by row)."""
        shapes, kinds = layout if layout is not None else ([d.shape for d in dims], [d.type for d in dims])
        width = len(dims)
        if any(shapes) or any(len(p) != width for p in points):    # (a non-scalar shape is a non-empty tuple)
            return None
        by_dim = list(zip(*points)) if n else [()] * len(dims)     # the transpose, done in C
        cols = [None] * len(dims)
        numeric = [j for j, kind in enumerate(kinds) if kind != 'categorical']
        if numeric:
            # ONE conversion for all numeric dimensions.  A str among the values makes numpy pick a string
            # dtype, None an object one, a list a ragged error: the general loop then raises the reference's
            # TypeError or rejects the row
            try:
                block = numpy.asarray([by_dim[j] for j in numeric]) if n else numpy.empty((len(numeric), 0))
            except ValueError:
                return None
            if block.dtype.kind not in 'fiub' or block.shape != (len(numeric), n):
                return None
            block = block.astype(numpy.float64, copy=False)
            for row, j in enumerate(numeric):
                cols[j] = block[row]
        for j, (dim, kind) in enumerate(zip(dims, kinds)):
            if kind == 'categorical':
                lookup = _category_lookup(dim)
                if lookup is None:
                    return None
                try:
                    cols[j] = numpy.fromiter(map(lookup.get, by_dim[j], _MINUS_ONES), dtype=numpy.int32, count=n)
                except TypeError:                                    # an unhashable value (list, ndarray)
                    return None
        return cols

    def columns_from_rows(self, points):
        """List of tuples -> (host columns, row_ok) where row_ok[i] is False
        when row i has the wrong arity or a mis-shaped value.  Integer values
        travel as float64 so the device can check integrality."""
        dims = list(self.values())
        n = len(points)
        fast = Space._columns_of_scalar_rows(dims, points, n, _layout_of(self))     # (self may be a reference Space)
        if fast is not None:
            return fast, numpy.ones(n, dtype=bool)
        widths = [int(numpy.prod(d.shape, dtype=numpy.int64)) if d.shape else 1 for d in dims]
        row_ok = numpy.ones(n, dtype=bool)
        cols = [numpy.zeros(n, dtype=numpy.int32 if d.type == 'categorical' else numpy.float64)
                for d, w in zip(dims, widths) for _ in range(w)]
        for i, point in enumerate(points):
            s = 0
            # zip() semantics of the reference: extra/missing components are ignored
            for j, dim in enumerate(dims):
                if j >= len(point):
                    break
                vals = _column_values(dim, point[j])
                if vals is None:
                    row_ok[i] = False
                else:
                    for e, v in enumerate(vals):
                        cols[s + e][i] = v
                s += widths[j]
            if len(point) < len(dims):
                # unchecked dimensions: fill with an always-valid value
                for j in range(len(point), len(dims)):
                    off = sum(widths[:j])
                    low, high = compile_space(self).dim_bounds[j]
                    fill = low if numpy.isfinite(low) else (high - 1 if numpy.isfinite(high) else 0)
                    for e in range(widths[j]):
                        cols[off + e][i] = fill
        return cols, row_ok

    def contains_points(self, points):
        """Batched ``point in space`` for a list of tuples -> list of bool."""
        if not self:
            return [False] * len(points)
        cols, row_ok = self.columns_from_rows(points)
        compiled = self.compiled()
        if len(points) <= HOST_SAMPLE_MAX:          # one library call: copy in, kernel, copy out, sync
            mask = compiled.contains_host(cols, len(points), int_as_f64=True)
        else:
            batch = ColumnBatch.from_numpy(compiled, cols, int_as_f64=True)
            mask = compiled.contains(batch, int_as_f64=True).cpu().numpy().astype(bool)
        return list(mask & row_ok)

    def interval(self, alpha=1.0):
        return [dim.categories if dim.type == 'categorical' else dim.interval(alpha)
                for dim in self.values()]

    def __getitem__(self, key):
        if isinstance(key, str):
            return super(Space, self).__getitem__(key)
        return list(self.values())[key]

    def __setitem__(self, key, value):
        if not isinstance(key, str):
            raise TypeError("Keys registered to Space must be string types. Provided: {}".format(key))
        if not isinstance(value, Dimension):
            raise TypeError("Values registered to Space must be Dimension types. Provided: {}".format(value))
        if key in self:
            raise ValueError("There is already a Dimension registered with this name. "
                             "Register it with another name. Provided: {}".format(key))
        self.__dict__.pop('_ob2_compiled', None)      # descriptor cache is stale now
        super(Space, self).__setitem__(key, value)

    def __contains__(self, value):
        """Dimension name registered? / point within the bounds of the space?"""
        if isinstance(value, str):
            return super(Space, self).__contains__(value)
        try:
            len(value)
        except TypeError as exc:
            raise TypeError("Can check only for dimension names or "
                            "for tuples with parameter values.") from exc
        if not self:
            return False
        return bool(self.contains_points([tuple(value)])[0])

    def __repr__(self):
        return "Space([{}])".format(',\n       '.join(map(str, self.values())))
