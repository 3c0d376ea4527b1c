"""Seeded list-returning calls return the REFERENCE's values (orion_b200/mt_stream.py over
``ob2_sample_from_base``): the unmodified reference (``oracle/_ref``, travels to the GPU box) and the
mirror are given the same seeds and must hand back the same points and leave the RandomState in the
same place -- src/orion/algo/space.py:110-136, 255-274, 281-294, 394-404, 456-484 and, for the
algorithm, src/orion/algo/random.py:24-33.  Bit for bit for uniform / norm / Integer / Categorical;
reciprocal goes through the device's ``exp`` and is held to 2 ulp.
tests/test_mt_stream.py checks the host logic alone, with a numpy stand-in for the kernel.
"""
from collections import OrderedDict

import numpy as np
import pytest
import torch

from oracle import ref_harness
from tests.test_mt_stream import CASES, FURTHER, make, random_dims, same_state, spaces

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_harness.available(), reason="no reference package staged (oracle/make_ref.py)")]

RECIPROCAL_RTOL = 4.5e-16        # 2 ulp


@pytest.fixture(autouse=True)
def reference():
    ref_harness.load()
    yield


def ref_classes():
    import orion.algo.space as ref
    return ref


def mirror():
    from orion_b200 import space
    return space


def equal_values(got, expect, exact=True, rtol=RECIPROCAL_RTOL, atol=0):
    g, e = np.asarray(got), np.asarray(expect)
    if g.shape != e.shape:
        return False
    if g.dtype == object or e.dtype == object:
        return bool(np.all(g == e))
    if exact:
        return np.array_equal(g, e)
    return bool(np.allclose(g, e, rtol=rtol, atol=atol))


@pytest.mark.parametrize('kind,args,kwargs', CASES)
@pytest.mark.parametrize('n', [1, 7, 40])
def test_dimension_sample_is_the_reference_s(kind, args, kwargs, n):
    ref_dim, dim = make(ref_classes(), kind, args, kwargs), make(mirror(), kind, args, kwargs)
    exact = args[0] in ('uniform', 'norm')
    # reciprocal: the device's exp, 2 ulp; a further prior: the device's inverse CDF, 1e-6 (an Integer over it: +-1)
    tol = dict(rtol=1e-6, atol=1 if kind == 'integer' else 0) if args[0] in FURTHER else dict(rtol=RECIPROCAL_RTOL)
    for seed in (3, 11):
        ref_rng, rng = np.random.RandomState(seed), np.random.RandomState(seed)
        for _call in range(2):
            try:
                expect = ref_dim.sample(n, ref_rng)
            except ValueError as exc:
                with pytest.raises(ValueError) as mine:
                    dim.sample(n, rng)
                assert str(mine.value) == str(exc)
                assert same_state(rng, ref_rng)
                continue
            got = dim.sample(n, rng)
            assert len(got) == len(expect) == n
            assert all(equal_values(g, e, exact, **tol) for g, e in zip(got, expect))
            assert all(type(g) is type(e) or np.asarray(g).dtype == np.asarray(e).dtype for g, e in zip(got, expect))
            assert same_state(rng, ref_rng)
        try:
            expect = ref_dim.sample(3, seed)        # an int: scipy seeds afresh for every value
        except ValueError:
            with pytest.raises(ValueError):
                dim.sample(3, seed)
        else:
            assert all(equal_values(g, e, exact, **tol) for g, e in zip(dim.sample(3, seed), expect))


def test_categorical_sample_is_the_reference_s():
    ref, mir = ref_classes(), mirror()
    for cats, kwargs in ((('asdfa', 2), {}),
                         (OrderedDict(zip(('asdfa', 2, 3, 4), (0.1, 0.2, 0.3, 0.4))), {}),
                         (('a', 'b', 'c'), {'shape': (2, 2)})):
        ref_dim, dim = ref.Categorical('yolo', cats, **kwargs), mir.Categorical('yolo', cats, **kwargs)
        assert all(equal_values(g, e) for g, e in zip(dim.sample(9, 5), ref_dim.sample(9, 5)))
        ref_rng, rng = np.random.RandomState(5), np.random.RandomState(5)
        assert all(equal_values(g, e) for g, e in zip(dim.sample(9, rng), ref_dim.sample(9, ref_rng)))
        assert same_state(rng, ref_rng)
        np.random.seed(9)
        expect = ref_dim.sample(9)
        np.random.seed(9)
        assert all(equal_values(g, e) for g, e in zip(dim.sample(9), expect))


@pytest.mark.parametrize('which', [0, 1])
@pytest.mark.parametrize('n', [1, 5, 64, 5000])
def test_space_sample_is_the_reference_s(which, n):
    ref_space, space = spaces(ref_classes())[which], spaces(mirror())[which]
    for seed in (1, 4):
        ref_rng, rng = np.random.RandomState(seed), np.random.RandomState(seed)
        for _call in range(2):
            try:
                expect = ref_space.sample(n, ref_rng)
            except ValueError as exc:
                with pytest.raises(ValueError) as mine:
                    space.sample(n, rng)
                assert str(mine.value) == str(exc)
                assert same_state(rng, ref_rng)
                continue
            got = space.sample(n, rng)
            assert len(got) == len(expect) == n
            for grow, erow in zip(got, expect):
                assert equal_values(grow[0], erow[0], exact=(which == 1))       # 'lr' is reciprocal in space 0
                assert all(equal_values(g, e) for g, e in zip(grow[1:], erow[1:]))
            assert same_state(rng, ref_rng)
        if n <= 64:                                 # an int seed repeats the call, whatever its outcome
            def outcome():
                try:
                    return [tuple(map(repr, p)) for p in space.sample(n, seed)]
                except ValueError as exc:
                    return str(exc)
            assert outcome() == outcome()


@pytest.mark.parametrize('case', range(20))
def test_random_spaces_sample_the_reference_s_points(case):
    """Generated spaces (tests/test_oracle_vs_reference_random.py::_random_dims) through the live reference
    and through the kernels; reciprocal values to 2 ulp (an Integer over one may then floor differently:
    those cases compare to +-1), everything else exact."""
    from tests import golden_util as gu
    dims_json = random_dims(case)
    ref_space, space = gu.product_space(dims_json, ref_classes()), gu.product_space(dims_json)
    for n in (1, 12):
        ref_rng, rng = np.random.RandomState(case), np.random.RandomState(case)
        try:
            expect = ref_space.sample(n, ref_rng)
        except ValueError as exc:
            with pytest.raises(ValueError) as mine:
                space.sample(n, rng)
            assert str(mine.value) == str(exc)
            assert same_state(rng, ref_rng)
            continue
        got = space.sample(n, rng)
        assert len(got) == len(expect) == n
        for grow, erow in zip(got, expect):
            for d, g, e in zip(dims_json, grow, erow):
                if d.get('prior') == 'reciprocal' and d['kind'] == 'integer':
                    assert np.all(np.abs(np.asarray(g) - np.asarray(e)) <= 1)
                else:
                    assert equal_values(g, e, exact=d.get('prior') != 'reciprocal'), (case, d)
        assert same_state(rng, ref_rng)


@pytest.mark.parametrize('which', [0, 1])
def test_sample_columns_on_the_reference_stream(which):
    """The tensor interface fed the reference's stream: the columns of the reference's seeded Space.sample
    (space 1 has user bounds: the rejections take the dimension-by-dimension path)."""
    ref_space, space = spaces(ref_classes())[which], spaces(mirror())[which]
    n = 3000
    ref_rng, rng = np.random.RandomState(8), np.random.RandomState(8)
    expect = ref_space.sample(n, ref_rng)
    batch = space.sample_columns(n, rng, reference_stream=True, with_fingerprint=True)
    assert batch.buf.is_cuda and batch.n == n and bool(batch.mask[:n].all().item()) and batch.fp is not None
    got = space.rows_from_columns(batch.to_numpy(), n)
    for grow, erow in zip(got, expect):
        assert equal_values(grow[0], erow[0], exact=(which == 1))
        assert all(equal_values(g, e) for g, e in zip(grow[1:], erow[1:]))
    assert same_state(rng, ref_rng)
    from orion_b200.space import Real, Space
    other = Space()
    other.register(Real('e', 'gamma', 2.0))
    with pytest.raises(NotImplementedError):
        other.sample_columns(4, 1, reference_stream=True)


def test_transformed_space_sample_is_transform_of_the_reference_s_points():
    from orion_b200.transformer import build_required_space
    ref_space, space = spaces(ref_classes())[1], spaces(mirror())[1]
    tspace = build_required_space('real', space)
    expect = tspace.transform_points(ref_space.sample(20, 7))
    got = tspace.sample(20, 7)
    for grow, erow in zip(got, expect):
        assert all(equal_values(g, e) for g, e in zip(grow, erow))


def test_gpurandom_suggests_what_the_reference_s_random_suggests():
    """Same numpy seed, same points: the reference's Random.suggest (random.py:24-33) against the plug-in
    over the SAME reference-built 64-dimension space, kernels doing the arithmetic."""
    from orion_b200 import compat
    compat.install()
    ref_space = ref_harness.build_space(ref_harness.MIXED64_ARGS)
    theirs = ref_harness.primary_algo(ref_space, 'random')
    mine = ref_harness.primary_algo(ref_space, 'gpurandom')
    for seed, num in ((1, 1), (2, 10), (3, 300)):
        np.random.seed(seed)
        expect = theirs.suggest(num)
        after = np.random.random_sample()
        np.random.seed(seed)
        got = mine.suggest(num)
        assert np.random.random_sample() == after
        assert len(got) == len(expect) == num
        for grow, erow in zip(got, expect):
            assert len(grow) == len(erow) == 64
            assert grow[:16] == erow[:16]                                        # uniform: bit for bit
            assert np.allclose(grow[16:24], erow[16:24], rtol=RECIPROCAL_RTOL, atol=0)      # loguniform
            assert grow[24:] == erow[24:]                                        # normal, integer, choices


def test_from_base_without_flags_is_from_uniform():
    """ob2_sample_from_base(flags = 0) and ob2_sample_from_uniform are the same launch."""
    from orion_b200 import _lib, spaces as product_spaces
    compiled = product_spaces.mixed64().compiled()
    n = 4096
    U = torch.rand((compiled.n_slots, n), dtype=torch.float64, device='cuda')
    a = compiled.sample_from_uniform(U, n)
    b = compiled.empty_columns(n)
    ok = torch.empty(n, dtype=torch.uint8, device='cuda')
    _lib.check(compiled.lib.ob2_sample_from_uniform(compiled.handle, n, n, U.data_ptr(), b.buf.data_ptr(), ok.data_ptr(),
                                                    None, compiled._stream()))
    assert torch.equal(a.buf, b.buf) and torch.equal(a.mask, ok)
    with pytest.raises(_lib.Ob2Error):
        _lib.check(compiled.lib.ob2_sample_from_base(compiled.handle, n, n, U.data_ptr(), 2, b.buf.data_ptr(),
                                                     ok.data_ptr(), None, compiled._stream()))
    # with the flag a norm slot returns variate * scale + loc exactly
    z = torch.randn((compiled.n_slots, n), dtype=torch.float64, device='cuda')
    mixed = torch.where(torch.tensor([d == _lib.DIST_NORM for d in compiled.dim_dists], device='cuda')[:, None], z, U)
    c = compiled.sample_from_uniform(mixed, n, normal_variates=True)
    cols_a, cols_c = a.to_numpy(), c.to_numpy()
    for s, dist in enumerate(compiled.dim_dists):
        if dist == _lib.DIST_NORM:
            assert np.array_equal(cols_c[s], z[s].cpu().numpy() * 1.0 + 0.0)
        else:
            assert np.array_equal(cols_c[s], cols_a[s])
