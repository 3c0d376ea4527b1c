"""The reference's own random stream for the list-returning interface.

``Space.sample(n, seed)`` / ``Dimension.sample(n, seed)`` of the reference draw every value through
numpy's MT19937 ``RandomState`` (reference ``src/orion/algo/space.py:133-136, 262-274, 294, 400-404,
482-484``): scipy's ``rvs`` asks it for ONE number per scalar -- ``random_sample()`` for ``uniform``
(``uniform_gen._rvs``), ``reciprocal`` / ``loguniform`` (the generic ``_ppf(uniform())``) and for
``RandomState.choice`` of a Categorical, ``standard_normal()`` for ``norm`` -- dimension by dimension,
sample by sample, drawing a rejected Real / Integer value again up to four times.

This module hands that stream to the device: the host only ASKS numpy for the numbers (that is where
the reference's random state lives; the caller may own it, or it is numpy's global one), the kernel
turns them into values (inverse CDF or scale/shift, ``floor``, cdf search, bounds check;
``ob2_sample_from_base``), and the accept / reject book-keeping of ``Real.sample`` is replayed on the
verdict bytes it returns.  A seeded call therefore returns what the reference returns for the same
seed and leaves the ``RandomState`` where the reference leaves it: bit for bit for ``uniform``,
``norm``, Integer and Categorical dimensions, to the last ulp or two of ``exp`` for ``reciprocal``.
The further priors scipy draws the same generic way (``fisk``, ``genpareto``, ``lomax``, ``invgamma``
... : no ``_rvs`` of their own) are fed uniforms too and come out within the device inverse CDF's
1e-6 (typically 1e-15) of the reference's values.

So are ``expon``, ``logistic`` and ``laplace``, whose numpy generators are that same inverse CDF of one uniform.

Priors whose scipy ``_rvs`` uses another numpy generator (``gamma`` -> ``standard_gamma``, ``beta``,
``poisson``, ``lognorm`` ...) and ``numpy.random.Generator`` seeds cannot be reproduced this way;
:func:`reproducible` says no for them and the callers keep the Philox stream of the tensor path.
"""
import numpy
import torch

from . import _lib
from .descriptor import round_up

UNIFORM, NORMAL, NOTHING = 'random_sample', 'standard_normal', 'nothing'

#: switch for the list-returning interface (``Space.sample``, ``Dimension.sample``,
#: ``Random.suggest`` without dedup): False keeps every draw on the Philox stream
ENABLED = True

_KIND_OF_DIST = {_lib.DIST_UNIFORM: UNIFORM, _lib.DIST_RECIPROCAL: UNIFORM, _lib.DIST_NORM: NORMAL,
                 _lib.DIST_CONST: NOTHING,       # scale == 0: rvs returns loc without drawing
                 None: UNIFORM}                  # Categorical: RandomState.choice -> random_sample


# priors with an ``_rvs`` of their own that still is ONE uniform per value put through the inverse CDF: numpy's
# legacy standard_exponential is -log(1 - u), logistic log(u / (1 - u)), laplace -log(2 - 2u) | log(2u)
# (numpy/random/src/legacy/legacy-distributions.c) -- checked against RandomState in tests/test_mt_stream.py
_UNIFORM_FED = (5, 7, 8)        # OB2_DIST_EXPON, OB2_DIST_LOGISTIC, OB2_DIST_LAPLACE


def stream_kinds(compiled):
    """Which RandomState method feeds each dimension of ``compiled``, or None when some prior's
    ``_rvs`` is not a function of ``random_sample`` / ``standard_normal`` numbers."""
    kinds = []
    for dist, by_inverse_cdf in zip(compiled.dim_dists, compiled.dim_generic_rvs):
        kind = _KIND_OF_DIST.get(dist)
        if kind is None and dist is not None and dist > _lib.DIST_NONE and (by_inverse_cdf or dist in _UNIFORM_FED):
            kind = UNIFORM                          # a further prior scipy draws as _ppf(uniform()): the device's _ppf
        if kind is None:
            return None
        kinds.append(kind)
    return kinds


def reproducible(compiled, rng):
    return ENABLED and isinstance(rng, numpy.random.RandomState) and stream_kinds(compiled) is not None


def _ask(rng, kind, count):
    if kind == UNIFORM:
        return rng.random_sample(count)
    if kind == NORMAL:
        return rng.standard_normal(count)
    return numpy.empty(0)


def _evaluate_on_device(compiled, base, n):
    """base: float64 [n_slots][n] -> (device ColumnBatch, ok[n] on the host) through the kernel."""
    ns = round_up(max(n, 1))
    block = numpy.full((compiled.n_slots, ns), 0.5)
    block[:, :n] = base
    batch = compiled.sample_from_uniform(torch.from_numpy(block).to(compiled.device), n, normal_variates=True)
    return batch, batch.mask[:n].cpu().numpy().astype(bool)


def _evaluate(compiled, base, n):
    """base: float64 [n_slots][n] -> (host columns, ok[n]) through the kernel."""
    if n <= _lib.HOST_SAMPLE_MAX:                   # numbers in, kernel, columns + verdicts out: one library call
        return compiled.sample_from_base_host(base, n)
    batch, ok = _evaluate_on_device(compiled, base, n)
    return batch.to_numpy(), ok


def _improbable(dim):
    return ValueError("Improbable bounds: (low={0}, high={1}). "
                      "Please make interval larger.".format(dim._low, dim._high))


def sample_space(dims, compiled, n, rng, on_device=False):
    """All dimensions in ONE launch, for the (overwhelmingly common) call in which no value is
    rejected: dimension-major stream, exactly ``[dim.sample(n, rng) for dim in space.values()]``.
    Returns the host columns (``on_device``: the device ColumnBatch), or None -- with ``rng`` put
    back where it was -- when some row missed its bounds and the caller has to go dimension by
    dimension."""
    kinds = stream_kinds(compiled)
    state = rng.get_state()
    base = numpy.full((compiled.n_slots, n), 0.5)
    firsts, widths = compiled.dim_first, compiled.dim_widths
    d = 0
    while d < len(kinds):
        # consecutive dimensions fed by the same RandomState method: ONE request for all their numbers
        # (the stream is dimension-major, so this is the sequence the separate requests would return)
        end = d + 1
        while end < len(kinds) and kinds[end] == kinds[d]:
            end += 1
        if kinds[d] != NOTHING:
            slots = sum(widths[d:end])
            numbers = _ask(rng, kinds[d], n * slots)
            if slots == end - d:                    # all scalar: dimension j's n numbers are row j
                base[firsts[d]:firsts[d] + slots] = numbers.reshape(slots, n)
            else:
                pos = 0
                for first, width in zip(firsts[d:end], widths[d:end]):
                    base[first:first + width] = numbers[pos:pos + n * width].reshape(n, width).T
                    pos += n * width
        d = end
    cols, ok = (_evaluate_on_device if on_device else _evaluate)(compiled, base, n)
    rejecting = [d.type in ('real', 'integer') for d in dims]
    if ok.all() or not any(rejecting):
        return cols
    rng.set_state(state)
    return None


def space_columns(dims, compiled, solo, n, rng):
    """Host columns of ``[dim.sample(n, rng) for dim in dims]`` (space.py:482-484): one launch when
    nothing is rejected, else dimension by dimension (``solo(dim)``: that dimension compiled alone)."""
    cols = sample_space(dims, compiled, n, rng)
    if cols is None:
        cols = []
        for dim in dims:
            cols.extend(sample_dimension(dim, solo(dim), n, rng))
    return cols


def space_batch(dims, compiled, solo, n, rng):
    """The same as a device ColumnBatch (``Space.sample_columns(..., reference_stream=True)``): the values of
    the reference's seeded ``Space.sample`` in the tensor interface.  The host still asks numpy for every
    number (a few ns each), so this is for reproducing a reference run, not for throughput."""
    from .columns import ColumnBatch
    batch = sample_space(dims, compiled, n, rng, on_device=True)
    if batch is None:
        batch = ColumnBatch.from_numpy(compiled, space_columns(dims, compiled, solo, n, rng))
        batch.mask = torch.ones(batch.n_stride, dtype=torch.uint8, device=compiled.device)   # accepted rows only
    return batch


def sample_dimension(dim, compiled, n, seed):
    """``dim.sample(n, seed)`` of the reference as host columns (one per element of the shape).

    ``compiled``: the dimension compiled as a one-dimension space.  An attempt consumes exactly
    ``width`` numbers whether it is accepted or not, so the stream is a sequence of equal chunks:
    the kernel evaluates a run of chunks at once and the four-attempt rule of ``Real.sample``
    (space.py:262-274) is replayed over their verdicts; the RandomState is then left exactly as
    many chunks further as the reference consumes."""
    from scipy._lib._util import check_random_state
    kind = stream_kinds(compiled)[0]
    width = compiled.n_slots
    rejects = dim.type in ('real', 'integer')       # bare Dimension and Categorical take what comes
    if n <= 0:
        return [numpy.empty(0, dtype=dt) for _, dt in compiled.col_meta]
    if isinstance(seed, (int, numpy.integer)) and dim.type != 'categorical':
        # rvs(random_state=<int>) seeds a fresh RandomState for EVERY draw (space.py:128-131 warns
        # about it): n copies of one value, and a rejected one is rejected four times over
        cols, ok = _evaluate(compiled, _ask(numpy.random.RandomState(seed), kind, width).reshape(1, -1).T
                             if kind != NOTHING else numpy.full((width, 1), 0.5), 1)
        if rejects and not ok[0]:
            raise _improbable(dim)
        return [numpy.repeat(c[:1], n) for c in cols]
    rng = check_random_state(seed)
    if kind == NOTHING:
        cols, ok = _evaluate(compiled, numpy.full((width, 1), 0.5), 1)
        if rejects and not ok[0]:
            raise _improbable(dim)
        return [numpy.repeat(c[:1], n) for c in cols]
    state = rng.get_state()
    chunks = n
    while True:
        cols, ok = _evaluate(compiled, _ask(rng, kind, chunks * width).reshape(chunks, width).T, chunks)
        if not rejects or ok[:n].all():
            take, used, failed = slice(0, n), n, False
        else:
            take, used, failed = _replay_attempts(ok, n)
            if take is None:                        # ran out of evaluated chunks: look further ahead
                rng.set_state(state)
                chunks = 2 * chunks + 8
                continue
        if used != chunks:                          # leave the RandomState where the reference leaves it
            rng.set_state(state)
            _ask(rng, kind, used * width)
        if failed:
            raise _improbable(dim)
        return [c[take] for c in cols]


def _replay_attempts(ok, n):
    """Real.sample over the verdicts of consecutive attempts: (indices taken, chunks consumed,
    gave up?) or (None, ..) when ``ok`` is too short to decide."""
    take = []
    j = 0
    for _ in range(n):
        for _attempt in range(4):
            if j >= len(ok):
                return None, j, False
            j += 1
            if ok[j - 1]:
                take.append(j - 1)
                break
        else:
            return take, j, True
    return numpy.asarray(take, dtype=numpy.int64), j, False
