"""The REFERENCE's own objects driving the kernels (``oracle/_ref`` travels to
the GPU box; ``oracle/make_ref.py`` stages it).  INTEGRATION.md's claims, checked:

* a ``Space`` built by the reference's ``SpaceBuilder().build_from``
  (src/orion/core/io/space_builder.py:236-261) compiles as it is and draws the
  same bytes as the mirror space;
* the reference's ``PrimaryAlgo(ref_space, 'gpurandom')``
  (src/orion/core/worker/primary_algo.py:25-56) runs on the device and every
  point passes the REFERENCE's ``__contains__`` (src/orion/algo/space.py:521-545);
* points the reference itself samples pass ``ob2_contains``.
"""
import numpy as np
import pytest
import torch

from oracle import ref_harness

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_harness.available(), reason="no reference package staged (oracle/make_ref.py)")]


@pytest.fixture(scope='module')
def ref_mixed64():
    space = ref_harness.build_space(ref_harness.MIXED64_ARGS)
    assert type(space).__module__ == 'orion.algo.space' and len(space) == 64
    yield space


def test_reference_built_space_compiles_to_the_mirror_s_bytes(ref_mixed64):
    from orion_b200 import spaces
    from orion_b200.descriptor import compile_space
    ref_c = compile_space(ref_mixed64)
    mir_c = spaces.mixed64().compiled()
    # (the reference's DSL keeps the namespace slash: '/u0')
    assert [name.lstrip('/') for name in ref_c.dim_names] == mir_c.dim_names
    assert ref_c.row_bytes == mir_c.row_bytes == 448
    assert ref_c.dim_bounds == mir_c.dim_bounds                 # the reference's interval() == the mirror's
    n, seed = 100_000, 20240003
    a = ref_c.sample(n, seed, with_mask=True, with_fingerprint=True)
    b = mir_c.sample(n, seed, with_mask=True, with_fingerprint=True)
    assert torch.equal(a.buf, b.buf) and torch.equal(a.fp, b.fp) and torch.equal(a.mask, b.mask)
    # and under the 'real' requirement (transform chain over the reference's dimensions)
    ta = compile_space(ref_mixed64, 'real').suggest(n, seed)
    tb = compile_space(spaces.mixed64(), 'real').suggest(n, seed)
    for x, y in zip(ta, tb):
        assert torch.equal(x.buf, y.buf)


def test_reference_primary_algo_runs_gpurandom_and_its_own_contains_accepts_every_point(ref_mixed64):
    from orion_b200 import compat
    klass = compat.install()
    algo = ref_harness.primary_algo(ref_mixed64, 'gpurandom')    # the reference's wrapper class
    assert type(algo).__module__ == 'orion.core.worker.primary_algo'
    assert type(algo.algorithm) is klass
    points = algo.suggest(200)            # PrimaryAlgo asserts `point in space` itself (primary_algo.py:53-54)
    assert len(points) == 200
    for p in points:
        assert len(p) == 64
        assert p in ref_mixed64           # the REFERENCE's Space.__contains__
        assert all(isinstance(v, float) for v in p[:32])
        assert all(float(v).is_integer() and -100 <= v < 100 for v in p[32:48])
        assert all(v in range(8) for v in p[48:])
    assert len(set(points)) == 200
    # the tensor fast path of the same object
    batch = algo.algorithm.suggest_columns(10_000)
    assert bool(batch.mask[:10_000].all().item())


def test_points_sampled_by_the_reference_pass_the_device_contains(ref_mixed64):
    from orion_b200.descriptor import compile_space
    from orion_b200.space import Space as MirrorSpace
    pts = ref_mixed64.sample(60, seed=7)                        # the reference's own sampler (MT19937)
    compiled = compile_space(ref_mixed64)
    cols, row_ok = MirrorSpace.columns_from_rows(ref_mixed64, pts)
    assert row_ok.all()
    mask = compiled.contains_host(cols, len(pts)) if hasattr(compiled, 'contains_host') else None
    if mask is None:
        from orion_b200.columns import ColumnBatch
        typed = [np.asarray(c, dtype={'f8': np.float64, 'i8': np.int64, 'i4': np.int32}[dt])
                 for c, (_u, dt) in zip(cols, compiled.col_meta)]
        mask = compiled.contains(ColumnBatch.from_numpy(compiled, typed))[:len(pts)].cpu().numpy()
    assert np.asarray(mask).astype(bool).all()
    # a point pushed out of one dimension's interval is refused by both
    bad = list(pts[0])
    bad[0] = 5.0
    assert tuple(bad) not in ref_mixed64
    cols, _ = MirrorSpace.columns_from_rows(ref_mixed64, [tuple(bad)])
    from orion_b200.columns import ColumnBatch
    typed = [np.asarray(c, dtype={'f8': np.float64, 'i8': np.int64, 'i4': np.int32}[dt])
             for c, (_u, dt) in zip(cols, compiled.col_meta)]
    assert not bool(compiled.contains(ColumnBatch.from_numpy(compiled, typed))[0].item())
