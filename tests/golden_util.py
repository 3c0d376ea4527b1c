"""Load the committed golden fixtures (made by tests/golden/make_golden.py
from the unmodified reference) into oracle DimSpecs and column lists."""
import json
import os

import numpy as np

from oracle.space_oracle import DimSpec

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
CASES = ['mixed_scalar', 'shaped', 'bounded_reject', 'loc_scale_kwargs']


def specs_from_json(dims):
    specs = []
    for d in dims:
        shape = tuple(d.get('shape') or ())
        if d['kind'] == 'categorical':
            cats = tuple(tuple(c) if isinstance(c, list) else c for c in d['categories'])
            probs = tuple(d['probs']) if d.get('probs') else tuple(np.tile(1. / len(cats), len(cats)))
            specs.append(DimSpec(d['name'], 'categorical', shape=shape, categories=cats, probs=probs))
        else:
            specs.append(DimSpec(d['name'], d['kind'], prior=d['prior'], args=tuple(d['args']),
                                 kwargs=dict(d.get('kwargs', {})), shape=shape,
                                 low=d.get('low', -np.inf), high=d.get('high', np.inf)))
    return specs


def load(name):
    z = np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False)
    dims = json.loads(str(z['spec']))
    out = dict(dims=dims, specs=specs_from_json(dims), n=int(z['n']), stream=z['stream'])
    if 'error' in z.files:
        out['error'] = str(z['error'])
        return out
    out['consumed'] = int(z['consumed'])
    out['cols'] = [z[k] for k in sorted(f for f in z.files if f.startswith('col'))]
    out['probes'] = [z[k] for k in sorted(f for f in z.files if f.startswith('probe0') or f.startswith('probe1'))]
    out['probe_verdict'] = z['probe_verdict']
    return out


def product_space(dims, module=None):
    """The same fixture spec as an orion_b200 Space (imports the product lazily) -- or, given the
    reference's ``orion.algo.space`` as ``module``, as a reference Space."""
    from collections import OrderedDict
    if module is None:
        from orion_b200 import space as module
    Categorical, Integer, Real, Space = module.Categorical, module.Integer, module.Real, module.Space
    space = Space()
    for d in dims:
        kw = dict(d.get('kwargs', {}))
        if d.get('shape'):
            kw['shape'] = tuple(d['shape'])
        if d['kind'] == 'categorical':
            cats = [tuple(c) if isinstance(c, list) else c for c in d['categories']]
            arg = OrderedDict(zip(cats, d['probs'])) if d.get('probs') else cats
            space.register(Categorical(d['name'], arg, **kw))
            continue
        if 'low' in d:
            kw['low'] = d['low']
        if 'high' in d:
            kw['high'] = d['high']
        klass = Real if d['kind'] == 'real' else Integer
        space.register(klass(d['name'], d['prior'], *d['args'], **kw))
    return space


def stream_to_U(specs, n, stream):
    """Dimension-major reference stream (rejection-free) -> U[S][n]."""
    U = []
    pos = 0
    for spec in specs:
        k = spec.n_slots
        U.extend(stream[pos:pos + n * k].reshape(n, k).T)
        pos += n * k
    return np.stack(U)
