"""The oracle against the reference's own outputs (committed fixtures) and the
known answers the reference's tests hold (tests/unittests/algo/test_space.py).
CPU only."""
import numpy as np
import pytest

from oracle import space_oracle as so
from tests import golden_util as gu


def test_philox_known_answers():
    """Random123 kat_vectors for philox4x32-10."""
    kat = [
        ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
        ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
        ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
         (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
    ]
    for ctr, key, want in kat:
        got = so.philox4x32_10(*ctr, *key)
        assert tuple(int(x) for x in got) == want


def test_u53_matches_numpy_random_sample():
    """u53 is MT19937's genrand_res53 construction: feed it MT's raw words."""
    rs = np.random.RandomState(5)
    st = np.random.RandomState(5)
    words = st.randint(0, 2 ** 32, size=8, dtype=np.uint64).astype(np.uint32)
    # randint(…, dtype=uint64 range 2**32) draws one 32-bit word per value
    want = rs.random_sample(4)
    got = so.u53(words[0::2], words[1::2])
    assert np.array_equal(got, want)


@pytest.mark.parametrize('name', gu.CASES)
def test_replay_matches_reference(name):
    """Same stream in -> bit-identical columns and the same number of uniforms
    consumed (i.e. the same rejections) as the reference."""
    g = gu.load(name)
    cols, consumed = so.replay_reference_stream(g['specs'], g['n'], g['stream'])
    assert consumed == g['consumed']
    assert len(cols) == len(g['cols'])
    for s, (got, want) in enumerate(zip(cols, g['cols'])):
        assert got.dtype == want.dtype, s
        assert np.array_equal(got, want), (name, s)


def test_improbable_bounds_raises_like_reference():
    g = gu.load('improbable')
    with pytest.raises(ValueError) as exc:
        so.replay_reference_stream(g['specs'], g['n'], g['stream'])
    assert 'Improbable bounds' in str(exc.value)
    assert str(exc.value) == g['error']


@pytest.mark.parametrize('name', ['mixed_scalar', 'shaped', 'loc_scale_kwargs'])
def test_columnar_from_uniform_matches_reference(name):
    """Rejection-free streams: U[S][N] laid out dimension-major is exactly the
    reference's stream, so the vectorised single-attempt path must
    reproduce the reference."""
    g = gu.load(name)
    n = g['n']
    U = []
    pos = 0
    for spec in g['specs']:
        k = spec.n_slots
        block = g['stream'][pos:pos + n * k].reshape(n, k).T
        U.extend(block)
        pos += n * k
    cols, ok = so.sample_from_uniform(g['specs'], np.stack(U))
    assert ok.all()
    for got, want in zip(cols, g['cols']):
        assert np.array_equal(got, want)


@pytest.mark.parametrize('name', gu.CASES)
def test_contains_matches_reference(name):
    g = gu.load(name)
    got = so.contains(g['specs'], g['probes'])
    assert np.array_equal(got, g['probe_verdict'])


def test_reference_unit_test_known_answers():
    """test_space.py:195-198 (Integer uniform(-3,6) seed 5 -> -2), :246-250
    ('asdfa' = index 0), :266-271 (probs .1/.2/.3/.4 seed 5 -> 2 = index 1)."""
    z = np.load(gu.GOLDEN + '/mt_seed5.npz')
    u = z['u']
    assert np.array_equal(np.random.RandomState(5).random_sample(4), u)
    i = so._draw(so.DimSpec('yolo', 'integer', prior='uniform', args=(-3, 6)), u[:1])
    assert int(i[0]) == -2 == int(z['integer_uniform_m3_6'])
    c = so._draw(so.DimSpec('yolo', 'categorical', categories=('asdfa', 2), probs=(0.5, 0.5)), u[:1])
    assert int(c[0]) == 0 == int(z['cat_pair_idx'])
    c = so._draw(so.DimSpec('yolo', 'categorical', categories=('asdfa', 2, 3, 4),
                            probs=(0.1, 0.2, 0.3, 0.4)), u[:1])
    assert int(c[0]) == 1 == int(z['cat_probs_idx'])


def test_interval_known_answers():
    """test_space.py:178-187, 200-201, 98."""
    S = so.DimSpec
    assert so.dim_interval(S('y', 'real', prior='norm', args=(0, 3), low=-3, high=3)) == (-3, 3)
    assert so.dim_interval(S('y', 'real', prior='alpha', args=(0.9,), low=-3, high=3)) == (0, 3)
    assert so.dim_interval(S('y', 'real', prior='uniform', args=(-2, 4), low=-3, high=3)) == (-2.0, 2.0)
    assert so.dim_interval(S('y', 'integer', prior='uniform', args=(-3, 6))) == (-3, 3)
    assert so.dim_interval(S('y', 'integer', prior='poisson', args=(5,)))[1] == np.inf


def test_categorical_frequencies():
    """test_space.py:283-296: law of large numbers on the category pmf."""
    spec = so.DimSpec('y', 'categorical', categories=('a', 'b', 'c', 'd'), probs=(0.1, 0.2, 0.3, 0.4))
    u = so.philox_uniform(1234, np.arange(200000), 0)
    idx = so._draw(spec, u)
    freq = np.bincount(idx, minlength=4) / len(idx)
    assert np.all(np.abs(freq - np.array(spec.probs)) < 0.01)
