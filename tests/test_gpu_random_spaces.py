"""The CUDA path against the oracle on randomly generated spaces (shaped
dimensions, user bounds with rejections, weighted categoricals): sample
(Philox), fused and unfused transform for every requirement, reverse,
contains, fingerprints."""
import numpy as np
import pytest
import torch

from oracle import space_oracle as so
from tests import golden_util as gu
from tests.test_oracle_vs_reference_random import _random_dims

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('case', range(20))
def test_random_space_matches_oracle(case):
    from orion_b200.transformer import build_required_space
    rng = np.random.RandomState(5000 + case)
    dims = _random_dims(rng)
    specs = gu.specs_from_json(dims)
    space = gu.product_space(dims)
    n = int(rng.choice([1, 7, 64, 513, 2000]))
    seed = int(rng.randint(1, 2 ** 31))
    base = int(rng.randint(0, 2 ** 40))
    requirement = [None, 'real', 'integer', 'numerical'][case % 4]
    compiled = build_required_space(requirement, space).compiled()
    tb = compiled.empty_columns(n, transformed=True)
    batch = compiled.sample(n, seed, base_index=base, with_fingerprint=True, transformed_out=tb)
    got = batch.to_numpy()
    want, failed = so.sample_philox(specs, n, seed, base_index=base)
    table = so.slot_table(specs)
    for s, (g, w) in enumerate(zip(got, want)):
        spec = specs[table[s][0]]
        assert g.dtype == w.dtype
        if spec.kind == 'real' and spec.prior != 'uniform':
            np.testing.assert_allclose(g, w, rtol=1e-6, atol=0)
        elif spec.kind == 'integer' and spec.prior != 'uniform':
            assert (g != w).sum() <= 1          # floor of a transcendental draw: ulp-level flips only
        else:
            assert np.array_equal(g, w), (case, s)
    assert np.array_equal(batch.mask[:n].cpu().numpy().astype(bool), ~failed)
    assert np.array_equal(compiled.contains(batch).cpu().numpy().astype(bool), so.contains(specs, got))
    lo, hi = batch.fingerprints_numpy()
    wlo, whi = so.fingerprint(got)
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)
    twant = so.transform(specs, requirement, got)
    for a, b in zip(tb.to_numpy(), twant):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    for a, b in zip(compiled.transform(batch).to_numpy(), twant):
        assert np.array_equal(a, b)
    bwant = so.reverse(specs, requirement, twant)
    for a, b in zip(compiled.reverse(tb).to_numpy(), bwant):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    # the one-kernel suggest over the same counters: every buffer as the staged kernels and the oracle have it
    sb, st, sk = compiled.suggest(n, seed, base_index=base)
    for a, b in zip(sb.to_numpy(), got):
        assert np.array_equal(a, b)
    for a, b in zip(st.to_numpy(), twant):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    for a, b in zip(sk.to_numpy(), bwant):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    assert torch.equal(sb.mask[:n], batch.mask[:n]) and torch.equal(sb.n_failed, batch.n_failed)
    slo, shi = sk.fingerprints_numpy()
    blo, bhi = so.fingerprint(bwant)
    assert np.array_equal(slo, blo) and np.array_equal(shi, bhi)
    # ... and the wire hand-over of a random subset is lossless
    keep_host = (rng.random_sample(n) < 0.6).astype(np.uint8)
    keep = torch.from_numpy(keep_host).cuda()
    in_space = batch.mask[:n].cpu().numpy().astype(bool)
    wire = batch.compact(keep & batch.mask[:n], wire=True)
    for a, b in zip(wire.to_numpy(), got):
        assert np.array_equal(a, b[keep_host.astype(bool) & in_space])
