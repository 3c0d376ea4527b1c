"""ob2_reverse_fingerprint: the reverse kernel also fingerprints the rows it
writes.  Lossless requirements: identical to the sampler's fingerprints;
'integer' (Quantize is lossy): the fingerprint of the reversed rows."""
import numpy as np
import pytest
import torch

from oracle import space_oracle as so
from tests import golden_util as gu

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('requirement', ['real', 'numerical', None, 'integer'])
@pytest.mark.parametrize('n', [1, 513, 20001])
def test_reverse_fingerprint(requirement, n):
    from orion_b200.transformer import build_required_space
    g = gu.load('mixed_scalar')
    dims = g['dims'] + [dict(name='sh', kind='categorical', categories=[1, 2, 3], shape=[2]),
                        dict(name='b2', kind='categorical', categories=['x', 'y'])]
    space = gu.product_space(dims)
    compiled = build_required_space(requirement, space).compiled()
    tb = compiled.empty_columns(n, transformed=True)
    batch = compiled.sample(n, 77, with_fingerprint=True, transformed_out=tb)
    plain = compiled.reverse(tb)
    back = compiled.reverse(tb, with_fingerprint=True)
    for a, b in zip(back.to_numpy(), plain.to_numpy()):
        assert np.array_equal(a, b)
    lo, hi = back.fingerprints_numpy()
    wlo, whi = so.fingerprint(back.to_numpy())                  # oracle fingerprint of the rows written
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)
    if requirement != 'integer':
        assert torch.equal(back.fp[:n], batch.fp[:n])          # == the sampler's fingerprints
    if n % 2:
        assert int(back.fp[n, 0].item()) == 0                   # padding entry of the odd tail


def test_pipeline_modes_agree():
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    from orion_b200.pipeline import CandidatePipeline
    from orion_b200.transformer import build_required_space
    compiled = build_required_space('real', spaces.mixed64()).compiled()
    n = 50_000
    rng = np.random.RandomState(3)
    index = [torch.from_numpy(rng.randint(0, 2 * n, size=n).astype(np.int64)).cuda() for _ in range(4)]
    kept = {}
    for mode in ('reverse', 'sample'):
        pipe = CandidatePipeline(compiled, TrialHistory(compiled, capacity=8 * n), n, fingerprint_in=mode)
        out = []
        for k in range(4):
            slot = pipe.submit(5, index=index[k])
            pipe.drain()
            out.append(slot.batch.keep[:n].clone())
        torch.cuda.synchronize()
        kept[mode] = out
    for a, b in zip(kept['reverse'], kept['sample']):
        assert torch.equal(a, b)
