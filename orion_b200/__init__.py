"""orion_b200 -- B200-native candidate generation for Orion.

One hot path, behind the reference's own Python plugin interface:
``Random.suggest`` -> ``Space.sample`` -> ``TransformedSpace`` transform/reverse
-> ``Space.__contains__`` -> trial-hash dedup, as hand-written sm_100a kernels
in ``libob2.so`` (C ABI: ``include/ob2.h``).  Importing the package is cheap;
the shared library is loaded (and its absence reported) on first compute use.
"""
__version__ = '0.1.0'

__all__ = ['Space', 'Dimension', 'Real', 'Integer', 'Categorical', 'build_required_space',
           'TransformedSpace', 'Random', 'PrimaryAlgo', 'TrialHistory', 'ShardedHistory']


def __getattr__(name):
    # lazy so that `import orion_b200` works where torch/CUDA are not wanted yet
    if name in ('Space', 'Dimension', 'Real', 'Integer', 'Categorical'):
        from . import space
        return getattr(space, name)
    if name in ('build_required_space', 'TransformedSpace'):
        from . import transformer
        return getattr(transformer, name)
    if name == 'Random':
        from .algo.random import Random
        return Random
    if name == 'PrimaryAlgo':
        from .primary_algo import PrimaryAlgo
        return PrimaryAlgo
    if name == 'TrialHistory':
        from .history import TrialHistory
        return TrialHistory
    if name == 'ShardedHistory':
        from .dist import ShardedHistory
        return ShardedHistory
    raise AttributeError(name)
