"""Algorithm interface and factory with the reference's contract.

Mirror of ``src/orion/algo/base.py`` (`BaseAlgorithm` :21-231,
`OptimizationAlgorithm` :235) and of the `Factory` discovery in
``src/orion/core/utils/__init__.py:46-116``: implementations are immediate
subclasses of `BaseAlgorithm`, advertised through the setuptools entry-point
group ``OptimizationAlgorithm`` and looked up by lower-cased class name.
"""
from abc import ABCMeta, abstractmethod
import logging

log = logging.getLogger(__name__)

ENTRY_POINT_GROUP = 'OptimizationAlgorithm'


class BaseAlgorithm(object, metaclass=ABCMeta):
    """What an optimisation algorithm can do: ``suggest`` points of a `Space`,
    ``observe`` their results.

    Keyword arguments given to ``__init__`` become the algorithm's tunable
    parameters (attributes, reported by ``configuration``).  A parameter whose
    value is ``{algo_name: {kwargs}}`` -- or the bare name of a registered
    algorithm -- is instantiated as a nested algorithm on the same space.
    """

    def __init__(self, space, **kwargs):
        log.debug("Creating Algorithm object of %s type with parameters:\n%s", type(self).__name__, kwargs)
        self._space = space
        self._param_names = list(kwargs.keys())
        OptimizationAlgorithm._discover()   # pylint:disable=protected-access
        for varname, param in kwargs.items():
            if isinstance(param, dict) and len(param) == 1:
                (sub_type, sub_kwargs), = param.items()
                if isinstance(sub_kwargs, dict):
                    param = OptimizationAlgorithm(sub_type, space, **sub_kwargs)
            elif isinstance(param, str) and param.lower() in OptimizationAlgorithm.typenames:
                param = OptimizationAlgorithm(param, space)
            setattr(self, varname, param)

    @abstractmethod
    def suggest(self, num=1):
        """Return ``num`` new points (tuples in space order)."""

    @abstractmethod
    def observe(self, points, results):
        """Take in evaluated ``points`` and their ``results`` (dicts with
        'objective', and optionally 'gradient', 'constraint')."""

    @property
    def is_done(self):
        return False

    def score(self, point):  # pylint:disable=no-self-use,unused-argument
        """Subjective expected performance of ``point``; no preference by default."""
        return 0

    def judge(self, point, measurements):  # pylint:disable=no-self-use,unused-argument
        """React to online ``measurements`` of a running trial; None by default."""
        return None

    @property
    def should_suspend(self):
        return False

    @property
    def configuration(self):
        """``{classname: {param: value}}``, nested algorithms expanded."""
        params = {}
        for attrname in self._param_names:
            if attrname.startswith('_'):
                continue
            attr = getattr(self, attrname)
            params[attrname] = attr.configuration if isinstance(attr, BaseAlgorithm) else attr
        return {type(self).__name__.lower(): params}

    @property
    def space(self):
        return self._space

    @space.setter
    def space(self, space_):
        self._space = space_
        for attr in self.__dict__.values():
            if isinstance(attr, BaseAlgorithm):
                attr.space = space_


class _AlgorithmFactory(ABCMeta):
    """``OptimizationAlgorithm(of_type, space, **kw)`` -> instance of the
    BaseAlgorithm subclass whose lower-cased name is ``of_type``."""

    _entry_points_loaded = False

    def _discover(cls):
        if not _AlgorithmFactory._entry_points_loaded:
            _AlgorithmFactory._entry_points_loaded = True
            try:
                from importlib.metadata import entry_points
                eps = entry_points()
                group = eps.select(group=ENTRY_POINT_GROUP) if hasattr(eps, 'select') else eps.get(ENTRY_POINT_GROUP, [])
                for ep in group:
                    try:
                        ep.load()
                    except ImportError as exc:  # e.g. a plugin for the reference package, which is not installed
                        log.debug("skipping %s plugin %s: %s", ENTRY_POINT_GROUP, ep.name, exc)
                    except Exception as exc:  # a broken third-party plugin must not take the factory down
                        log.warning("could not load %s plugin %s: %s", ENTRY_POINT_GROUP, ep.name, exc)
            except ImportError:
                pass
        found = [c for c in BaseAlgorithm.__subclasses__() if c.__name__ != cls.__name__]
        for c in found:
            if c not in cls.types:
                cls.types.append(c)
                cls.typenames.append(c.__name__.lower())

    def __init__(cls, name, bases, namespace):
        super(_AlgorithmFactory, cls).__init__(name, bases, namespace)
        cls.types = []
        cls.typenames = []

    def __call__(cls, of_type, *args, **kwargs):
        cls._discover()
        for klass in cls.types:
            if klass.__name__.lower() == of_type.lower():
                return klass(*args, **kwargs)
        raise NotImplementedError(
            "Could not find implementation of {0}, type = '{1}'\n"
            "Currently, there is an implementation for types:\n{2}".format(
                BaseAlgorithm.__name__, of_type, cls.typenames))


class OptimizationAlgorithm(BaseAlgorithm, metaclass=_AlgorithmFactory):  # pylint:disable=abstract-method
    """Factory entry: ``OptimizationAlgorithm('random', space)``."""
