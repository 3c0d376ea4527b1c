#!/usr/bin/env python
"""Generate golden vectors by running the UNMODIFIED reference under a replayed
uniform stream.  Runs only in the build container (needs /root/reference);
its outputs (``tests/golden/*.npz``) are committed and travel to the GPU box.

    python tests/golden/make_golden.py

Harness accommodations (none edits the reference):
  * ``numpy.object = object`` -- ``space.py:399,422`` use the alias numpy
    removed in 1.24.
  * ``InjectRS``: a ``numpy.random.RandomState`` subclass passed as ``seed``;
    ``check_random_state`` (space.py:482) passes instances through, so every
    ``uniform`` / ``random_sample`` the reference and scipy issue is served from
    our stream.  ``standard_normal`` (used by ``norm_gen._rvs``) is served as
    ``ndtri(next uniform)`` -- a harness convention, since the reference's own
    normal draw is not a function of one uniform.
"""
import json
import os
import sys
from collections import OrderedDict

import numpy
import numpy as np

numpy.object = object  # noqa: shim for the reference on numpy >= 1.24
sys.path.insert(0, '/root/reference/src')

from scipy.special import ndtri  # noqa: E402
from orion.algo.space import Categorical, Integer, Real, Space  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


class InjectRS(np.random.RandomState):
    """Replay a fixed stream of uniforms through the RandomState interface."""

    def __init__(self, stream):
        super().__init__(0)
        self.stream = np.asarray(stream, dtype=np.float64)
        self.pos = 0

    def _take(self, size):
        if size is None:
            k, shape = 1, None
        else:
            shape = tuple(np.atleast_1d(size).astype(int)) if not isinstance(size, tuple) else size
            k = int(np.prod(shape)) if len(shape) else 1
        out = self.stream[self.pos:self.pos + k]
        if len(out) < k:
            raise RuntimeError("stream exhausted")
        self.pos += k
        if size is None:
            return np.float64(out[0])
        return out.reshape(shape).copy()

    def random_sample(self, size=None):
        return self._take(size)

    def uniform(self, low=0.0, high=1.0, size=None):
        return low + (high - low) * self._take(size)

    def standard_normal(self, size=None):
        return ndtri(self._take(size))


def build(dims):
    """dims: list of spec dicts (the same JSON that is stored in the fixture)."""
    space = Space()
    for d in dims:
        kw = dict(d.get('kwargs', {}))
        if d.get('shape'):
            kw['shape'] = tuple(d['shape'])
        if d['kind'] == 'categorical':
            cats = [tuple(c) if isinstance(c, list) else c for c in d['categories']]
            if d.get('probs'):
                arg = OrderedDict(zip(cats, d['probs']))
            else:
                arg = cats
            space.register(Categorical(d['name'], arg, **kw))
            continue
        if 'low' in d:
            kw['low'] = d['low']
        if 'high' in d:
            kw['high'] = d['high']
        klass = Real if d['kind'] == 'real' else Integer
        space.register(klass(d['name'], d['prior'], *d['args'], **kw))
    return space


def columns(space, dims, points):
    """list of tuples -> one float64/int64/int32 array per scalar slot."""
    cols = []
    for j, d in enumerate(dims):
        vals = [np.asarray(p[j]) for p in points]
        if d['kind'] == 'categorical':
            cats = list(space[j].categories)
            idx = [np.vectorize(lambda v: cats.index(v), otypes='i')(np.asarray(p[j], dtype=object))
                   for p in points]
            block = np.stack([np.asarray(i).reshape(-1) for i in idx], axis=1).astype(np.int32)
        elif d['kind'] == 'integer':
            block = np.stack([v.reshape(-1) for v in vals], axis=1).astype(np.int64)
        else:
            block = np.stack([v.reshape(-1) for v in vals], axis=1).astype(np.float64)
        cols.extend(block)
    return cols


CASES = OrderedDict()

CASES['mixed_scalar'] = dict(n=64, seed=11, dims=[
    dict(name='u', kind='real', prior='uniform', args=[-1, 2]),
    dict(name='l', kind='real', prior='reciprocal', args=[1e-5, 1]),
    dict(name='g', kind='real', prior='norm', args=[0.9, 0.1]),
    dict(name='i', kind='integer', prior='uniform', args=[-100, 200]),
    dict(name='i2', kind='integer', prior='uniform', args=[-3, 6]),
    dict(name='li', kind='integer', prior='reciprocal', args=[1, 1000]),
    dict(name='c', kind='categorical', categories=[0, 1, 2, 3, 4, 5, 6, 7]),
    dict(name='cp', kind='categorical', categories=['asdfa', 2, 3, 4], probs=[0.1, 0.2, 0.3, 0.4]),
])

CASES['shaped'] = dict(n=16, seed=12, dims=[
    dict(name='yolo', kind='categorical', categories=['asdfa', 2, 3, 4],
         probs=[0.1, 0.2, 0.3, 0.4], shape=[2, 2]),
    dict(name='yolo2', kind='integer', prior='uniform', args=[-3, 6], shape=[3]),
    dict(name='yolo3', kind='real', prior='norm', args=[0.9]),
    dict(name='yolo4', kind='real', prior='uniform', args=[-3, 4], shape=[4, 4]),
])

CASES['bounded_reject'] = dict(n=40, seed=13, dims=[
    dict(name='a', kind='real', prior='norm', args=[0, 1], low=-1.5, high=1.5),
    dict(name='b', kind='real', prior='uniform', args=[-2, 4], low=-1.9, high=3.0),
    dict(name='c', kind='integer', prior='norm', args=[0, 3], low=-7, high=7, shape=[2]),
    dict(name='d', kind='real', prior='reciprocal', args=[1e-3, 10], low=1e-2),
])

CASES['improbable'] = dict(n=8, seed=14, expect_error=True, dims=[
    dict(name='yolo', kind='real', prior='norm', args=[0, 2], low=-2, high=2, shape=[4, 4]),
])

CASES['loc_scale_kwargs'] = dict(n=32, seed=15, dims=[
    dict(name='a', kind='real', prior='uniform', args=[], kwargs=dict(loc=3.0, scale=0.5)),
    dict(name='b', kind='real', prior='reciprocal', args=[0.5, 8.0], kwargs=dict(loc=-1.0, scale=2.0)),
    dict(name='c', kind='real', prior='norm', args=[], kwargs=dict(loc=-4.0, scale=2.5)),
    dict(name='d', kind='integer', prior='norm', args=[10, 4]),
])


def contains_probe(space, dims, points, rng):
    """Points around the accepted samples: perturbed, out of range, wrong kind."""
    probes = list(points)
    for p in points[:16]:
        q = list(p)
        j = int(rng.randint(len(q)))
        d = dims[j]
        if d['kind'] == 'categorical':
            continue
        v = np.asarray(q[j], dtype=np.float64)
        q[j] = v + rng.choice([-1e3, -0.5, 0.25, 0.5, 1.0, 1e3])
        if v.shape == ():
            q[j] = float(q[j])
        probes.append(tuple(q))
    verdict = [bool(p in space) for p in probes]
    return probes, verdict


def main():
    for name, case in CASES.items():
        dims = case['dims']
        space = build(dims)
        n = case['n']
        rng = np.random.RandomState(case['seed'])
        n_slots = sum(int(np.prod(d.get('shape', []) or [1])) for d in dims)
        stream = rng.random_sample(n * n_slots * 8 + 64)
        inj = InjectRS(stream)
        out = dict(spec=json.dumps(dims), n=n)
        try:
            points = space.sample(n, seed=inj)
        except ValueError as exc:
            assert case.get('expect_error'), exc
            out['error'] = str(exc)
            out['stream'] = stream[:inj.pos]
            np.savez_compressed(os.path.join(HERE, name + '.npz'), **out)
            print(name, 'raised as expected:', exc)
            continue
        assert not case.get('expect_error')
        cols = columns(space, dims, points)
        out['consumed'] = inj.pos
        out['stream'] = stream[:inj.pos]
        for s, c in enumerate(cols):
            out['col%03d' % s] = c
        probes, verdict = contains_probe(space, dims, points, rng)
        pcols_f = []
        for j, d in enumerate(dims):
            if d['kind'] == 'categorical':
                cats = list(space[j].categories)
                block = np.stack([np.vectorize(lambda v: cats.index(v), otypes='i')(
                    np.asarray(p[j], dtype=object)).reshape(-1) for p in probes], axis=1).astype(np.int32)
            else:
                # probes are stored as float64 for Real AND Integer so that the
                # mod(x, 1) == 0 rule of Integer.__contains__ is exercised
                block = np.stack([np.asarray(p[j], dtype=np.float64).reshape(-1) for p in probes], axis=1)
            pcols_f.extend(block)
        for s, c in enumerate(pcols_f):
            out['probe%03d' % s] = c
        out['probe_verdict'] = np.asarray(verdict, dtype=bool)
        np.savez_compressed(os.path.join(HERE, name + '.npz'), **out)
        print(name, 'n=%d slots=%d consumed=%d rejections=%d contains_true=%d/%d' % (
            n, n_slots, inj.pos, inj.pos - n * n_slots, sum(verdict), len(verdict)))

    # reference's own golden values (tests/unittests/algo/test_space.py:195-198,
    # 246-250, 266-271): MT19937-seeded draws -> which uniforms they consumed
    rs = np.random.RandomState(5)
    u = rs.random_sample(4)
    np.savez_compressed(os.path.join(HERE, 'mt_seed5.npz'), u=u,
                        integer_uniform_m3_6=np.int64(Integer('yolo', 'uniform', -3, 6).sample(seed=np.random.RandomState(5))[0]),
                        cat_probs_idx=np.int32(['asdfa', 2, 3, 4].index(
                            Categorical('yolo', OrderedDict(zip(('asdfa', 2, 3, 4), (0.1, 0.2, 0.3, 0.4)))).sample(
                                seed=np.random.RandomState(5))[0])),
                        cat_pair_idx=np.int32(('asdfa', 2).index(
                            Categorical('yolo', ('asdfa', 2)).sample(seed=np.random.RandomState(5))[0])))
    print('mt_seed5 written')


if __name__ == '__main__':
    main()
