"""Algorithm API mirror: BaseAlgorithm, the OptimizationAlgorithm factory, Random."""
from .base import BaseAlgorithm, OptimizationAlgorithm  # noqa: F401
from .random import Random  # noqa: F401
