"""Product Space -> oracle DimSpec list (reads only public/reference-named
attributes of the dimensions, so it also works on the reference's classes)."""
import numpy as np

from oracle.space_oracle import DimSpec


def specs_from_space(space):
    specs = []
    for dim in space.values():
        shape = tuple(dim.shape)
        if dim.type == 'categorical':
            specs.append(DimSpec(dim.name, 'categorical', shape=shape,
                                 categories=tuple(dim.categories), probs=tuple(dim._probs)))
        else:
            kind = 'real' if dim.type == 'dimension' else dim.type
            specs.append(DimSpec(dim.name, kind, prior=dim._prior_name, args=tuple(dim._args),
                                 kwargs=dict(dim._kwargs), shape=shape,
                                 low=getattr(dim, '_low', -np.inf), high=getattr(dim, '_high', np.inf)))
    return specs
