"""The sampler's draw plan as the oracle restates it (oracle/space_oracle.py: draw_plan,
uses_32_bit_uniform), and the two integer shortcuts the CUDA sampler takes for 32-bit
slots, proven equal to the floating-point statement on edge and random words.  CPU only;
the device side is held to the same oracle by tests/test_gpu_*.py."""
import numpy as np

from oracle import space_oracle as so
from oracle.space_oracle import DimSpec


def _mixed64():
    specs = [DimSpec('u%d' % i, 'real', prior='uniform', args=(-1, 2)) for i in range(16)]
    specs += [DimSpec('l%d' % i, 'real', prior='reciprocal', args=(1e-5, 1)) for i in range(8)]
    specs += [DimSpec('g%d' % i, 'real', prior='norm', args=(0, 1)) for i in range(8)]
    specs += [DimSpec('i%d' % i, 'integer', prior='uniform', args=(-100, 200)) for i in range(16)]
    specs += [DimSpec('c%d' % i, 'categorical', categories=tuple(range(8)), probs=tuple(np.tile(1. / 8, 8)))
              for i in range(16)]
    return specs


def test_philox_7_round_known_answers():
    """Random123's kat_vectors for philox4x32 with 7 rounds."""
    kat = [((0, 0, 0, 0), (0, 0), (0x5f6fb709, 0x0d893f64, 0x4f121f81, 0x4f730a48)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x5207ddc2, 0x45165e59, 0x4d8ee751, 0x8c52f662)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0x4dfccaba, 0x190a87f0, 0xc47362ba, 0xb6b5242a))]
    for ctr, key, want in kat:
        got = so.philox4x32_10(*ctr, *key, rounds=7)
        assert tuple(int(x) for x in got) == want


def test_plan_of_real_spaces_is_the_two_slots_per_block_layout():
    specs = [DimSpec('u%d' % i, 'real', prior='uniform', args=(-1, 2)) for i in range(5)]
    specs.append(DimSpec('v', 'real', prior='norm', args=(0, 1), shape=(3,)))
    plan = so.draw_plan(specs)
    assert plan == [(s // 2, 2 * (s % 2), 53) for s in range(8)]
    a = so.philox_uniform(9, np.arange(100), 3)
    b = so.philox_uniform(9, np.arange(100), plan[3])
    assert np.array_equal(a, b)


def test_plan_of_mixed64_packs_four_narrow_slots_per_block():
    plan = so.draw_plan(_mixed64())
    assert [p[2] for p in plan] == [53] * 32 + [32] * 32
    assert plan[31] == (15, 2, 53) and plan[32] == (16, 0, 32) and plan[35] == (16, 3, 32) and plan[36] == (17, 0, 32)
    assert plan[-1] == (23, 3, 32)                      # 24 blocks per candidate instead of 32


def test_plan_alignment_when_widths_alternate():
    specs = [DimSpec('c', 'categorical', categories=(0, 1, 2), probs=(0.2, 0.3, 0.5)),
             DimSpec('x', 'real', prior='uniform', args=(0, 1)),
             DimSpec('k', 'integer', prior='uniform', args=(0, 10)),
             DimSpec('d', 'categorical', categories=(0, 1), probs=(0.5, 0.5)),
             DimSpec('y', 'real', prior='uniform', args=(0, 1)),
             DimSpec('z', 'real', prior='uniform', args=(0, 1))]
    assert so.draw_plan(specs) == [(0, 0, 32), (0, 2, 53), (1, 0, 32), (1, 1, 32), (1, 2, 53), (2, 0, 53)]


def test_which_slots_take_32_bits():
    yes = [DimSpec('k', 'integer', prior='uniform', args=(-100, 200)),
           DimSpec('k', 'integer', prior='uniform', args=(0, 2 ** 20)),
           DimSpec('c', 'categorical', categories=(0, 1, 2), probs=(0.1, 0.2, 0.7)),
           DimSpec('c', 'categorical', categories=(0,), probs=(1.0,))]
    no = [DimSpec('k', 'integer', prior='uniform', args=(0, 2 ** 20 + 1)),
          DimSpec('k', 'integer', prior='uniform', args=(0.5, 10)),
          DimSpec('k', 'integer', prior='uniform', args=(0, 10), high=5),          # user bound inside the support
          DimSpec('k', 'integer', prior='norm', args=(0, 5)),
          DimSpec('x', 'real', prior='uniform', args=(0, 1)),
          DimSpec('c', 'categorical', categories=(0, 1), probs=(1.0 - 2.0 ** -40, 2.0 ** -40))]
    assert all(so.uses_32_bit_uniform(s) for s in yes)
    assert not any(so.uses_32_bit_uniform(s) for s in no)


def _edge_words(rng, extra=()):
    w = np.concatenate([rng.randint(0, 2 ** 32, size=200000, dtype=np.uint64),
                        np.array([0, 1, 2, 2 ** 31 - 1, 2 ** 31, 2 ** 31 + 1, 2 ** 32 - 2, 2 ** 32 - 1], dtype=np.uint64),
                        np.asarray(extra, dtype=np.uint64)])
    return w


def test_integer_shortcut_equals_floor_of_the_double_arithmetic():
    """loc + ((w * scale) >> 32) == floor((w / 2^32) * scale + loc), rounded multiply then rounded add."""
    rng = np.random.RandomState(3)
    for loc, scale in [(-100, 200), (0, 1), (-2 ** 20, 2 ** 20), (2 ** 20 - 7, 2 ** 20), (5, 999983), (-3, 6)]:
        # words around every integer boundary of w * scale / 2^32
        ks = rng.randint(0, scale, size=2000, dtype=np.int64)
        near = ((ks.astype(np.uint64) << np.uint64(32)) + np.uint64(scale - 1)) // np.uint64(scale)
        near = np.concatenate([near, near - 1, near + 1]) % np.uint64(2 ** 32)
        w = _edge_words(rng, near)
        u = w.astype(np.float64) * 2.0 ** -32
        want = np.floor(u * float(scale) + float(loc)).astype(np.int64)
        got = loc + ((w * np.uint64(scale)) >> np.uint64(32)).astype(np.int64)
        assert np.array_equal(got, want), (loc, scale)
        assert got.min() >= loc and got.max() < loc + scale


def test_categorical_threshold_count_equals_searchsorted():
    """#{j < k-1 : ceil(cdf[j] 2^32) <= w} == cdf.searchsorted(w / 2^32, side='right'); and with
    equal probabilities whose thresholds are exactly ceil((j+1) 2^32 / k) the count is (w * k) >> 32."""
    rng = np.random.RandomState(4)
    tables = [np.tile(1. / 8, 8), np.tile(1. / 100, 100), np.tile(1. / 3, 3), np.array([0.1, 0.2, 0.3, 0.4]),
              rng.dirichlet(np.ones(17)), np.array([1.0]), np.array([0.5, 0.5])]
    for probs in tables:
        cdf = so.categorical_cdf(probs)
        k = len(cdf)
        thr = np.ceil(cdf[:-1] * 4294967296.0)
        assert np.all(thr <= 4294967295.0)
        thr = thr.astype(np.uint64)
        w = _edge_words(rng, np.concatenate([thr, thr - np.uint64(1), thr + np.uint64(1)]) % np.uint64(2 ** 32))
        want = cdf.searchsorted(w.astype(np.float64) * 2.0 ** -32, side='right')
        got = (thr[None, :] <= w[:, None]).sum(axis=1) if k > 1 else np.zeros(len(w), dtype=np.int64)
        assert np.array_equal(got, want), k
        assert want.max() <= k - 1
        even = np.array([(((j + 1) << 32) + k - 1) // k for j in range(k - 1)], dtype=np.uint64)
        if np.array_equal(even, thr):
            assert np.array_equal((w * np.uint64(k)) >> np.uint64(32), want.astype(np.uint64)), k
    # 8 equal categories: the shortcut applies (1/8 is exact)
    cdf8 = so.categorical_cdf(np.tile(1. / 8, 8))
    assert np.array_equal(np.ceil(cdf8[:-1] * 2.0 ** 32), np.arange(1, 8) * 2.0 ** 29)


def test_sample_philox_uses_the_plan_and_rounds():
    specs = _mixed64()
    a, fa = so.sample_philox(specs, 2000, 11)
    b, fb = so.sample_philox(specs, 2000, 11, rounds=7)
    assert not fa.any() and not fb.any()
    assert not np.array_equal(a[0], b[0])                          # other rounds, other draws
    plan = so.draw_plan(specs)
    u = so.philox_uniform(11, np.arange(2000), plan[40])           # an Integer slot: one word
    assert np.array_equal(a[40], np.floor(u * 200.0 - 100.0).astype(np.int64))
    w = so.philox4x32_10(np.arange(2000, dtype=np.uint32), 0, plan[40][0], 0, 11, 0)[plan[40][1]]
    assert np.array_equal(a[40], -100 + ((w.astype(np.uint64) * np.uint64(200)) >> np.uint64(32)).astype(np.int64))
    assert all(0 <= c.min() and c.max() <= 7 for c in a[48:])
