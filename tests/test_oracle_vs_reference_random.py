"""The oracle against the LIVE reference on randomly generated spaces.

Runs only where the reference checkout exists (the build container); the
committed fixtures in tests/golden/ carry the same kind of evidence to the GPU
box.  Each case builds a random space, replays one random uniform stream
through the unmodified reference (`Space.sample(n, seed=InjectRS)`,
space.py:456-484) and through `replay_reference_stream`, and compares values,
stream consumption (= rejections) and `point in space` verdicts."""
import importlib.util
import os
import sys

import numpy as np
import pytest

REF = '/root/reference/src'
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present")


def _load_generator():
    here = os.path.dirname(os.path.abspath(__file__))
    spec = importlib.util.spec_from_file_location('make_golden', os.path.join(here, 'golden', 'make_golden.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)             # imports the reference (adds REF to sys.path, shims numpy.object)
    return mod


def _random_dims(rng):
    dims = []
    for j in range(rng.randint(1, 7)):
        kind = rng.choice(['real', 'integer', 'categorical'])
        shape = [None, None, [2], [3], [2, 2]][rng.randint(5)]
        d = dict(name='d%d' % j, kind=str(kind))
        if shape:
            d['shape'] = shape
        if kind == 'categorical':
            k = int(rng.randint(1, 7))
            d['categories'] = [int(c) if c % 2 else 'c%d' % c for c in range(k)]
            if rng.rand() < 0.5 and k > 1:
                p = rng.dirichlet(np.ones(k))
                p = np.round(p, 3)
                p[-1] = 1.0 - p[:-1].sum()
                if (p > 0).all():
                    d['probs'] = [float(x) for x in p]
        else:
            prior = str(rng.choice(['uniform', 'reciprocal', 'norm']))
            d['prior'] = prior
            if prior == 'uniform':
                lo = float(np.round(rng.uniform(-50, 50), 2))
                d['args'] = [lo, float(np.round(rng.uniform(0.5, 60), 2))]
            elif prior == 'reciprocal':
                a = float(10.0 ** rng.randint(-5, 1))
                d['args'] = [a, a * float(10.0 ** rng.randint(1, 5))]
            else:
                d['args'] = [float(np.round(rng.uniform(-5, 5), 2)), float(np.round(rng.uniform(0.1, 4), 2))]
                if rng.rand() < 0.5:      # user bounds around +-2.5 sigma: a few rejections, no failures
                    d['low'] = d['args'][0] - 2.5 * d['args'][1]
                    d['high'] = d['args'][0] + 2.5 * d['args'][1]
        dims.append(d)
    return dims


@pytest.mark.parametrize('case', range(25))
def test_random_space_matches_live_reference(case):
    from oracle import space_oracle as so
    from tests import golden_util as gu
    gen = _load_generator()
    try:
        rng = np.random.RandomState(1000 + case)
        dims = _random_dims(rng)
        space = gen.build(dims)
        specs = gu.specs_from_json(dims)
        n = 12
        n_slots = sum(s.n_slots for s in specs)
        stream = rng.random_sample(n * n_slots * 6 + 16)
        inj = gen.InjectRS(stream)
        try:
            points = space.sample(n, seed=inj)
        except ValueError as exc:            # "Improbable bounds": the restatement must fail the same way
            with pytest.raises(ValueError) as mine:
                so.replay_reference_stream(specs, n, stream)
            assert str(mine.value) == str(exc)
            return
        want = gen.columns(space, dims, points)
        got, consumed = so.replay_reference_stream(specs, n, stream)
        assert consumed == inj.pos
        for s, (g, w) in enumerate(zip(got, want)):
            assert g.dtype == w.dtype and np.array_equal(g, w), (case, s)
        # membership: the samples, and perturbed copies
        probes = list(points)
        for p in points[:6]:
            q = list(p)
            j = int(rng.randint(len(q)))
            if dims[j]['kind'] != 'categorical':
                q[j] = np.asarray(q[j], dtype=np.float64) + rng.choice([-1e3, -0.5, 0.5, 1e3])
                if not dims[j].get('shape'):
                    q[j] = float(q[j])
            probes.append(tuple(q))
        verdict = np.array([p in space for p in probes])
        pcols = []
        for j, d in enumerate(dims):
            if d['kind'] == 'categorical':
                cats = list(space[j].categories)
                block = np.stack([np.vectorize(lambda v: cats.index(v), otypes='i')(
                    np.asarray(p[j], dtype=object)).reshape(-1) for p in probes], axis=1).astype(np.int32)
            else:
                block = np.stack([np.asarray(p[j], dtype=np.float64).reshape(-1) for p in probes], axis=1)
            pcols.extend(block)
        assert np.array_equal(so.contains(specs, pcols), verdict)
    finally:
        if REF in sys.path:
            sys.path.remove(REF)
        for name in [m for m in sys.modules if m == 'orion' or m.startswith('orion.')]:
            del sys.modules[name]
