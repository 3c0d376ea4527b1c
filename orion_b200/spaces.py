"""Search-space construction from the reference's prior DSL, and the named
synthetic spaces of the benchmark.

``build_dimension(name, 'uniform(-1, 1)')`` accepts the expressions the
reference's ``DimensionBuilder`` documents (``src/orion/core/io/space_builder.py:69-215``):
``uniform(a, b)`` means [a, b) (scipy's loc/scale = a, b - a); ``loguniform`` is
scipy's ``reciprocal``; ``normal``/``gaussian`` is ``norm``; ``choices([...])``
or ``choices({cat: p})`` is Categorical; ``discrete=True`` selects Integer.
Expressions are parsed with ``ast`` (literals only) rather than ``eval``.
"""
import ast

from .space import Categorical, Integer, Real, Space


def _literal_call(expression):
    try:
        node = ast.parse(expression.strip(), mode='eval').body
    except SyntaxError:
        node = None
    if not isinstance(node, ast.Call) or not isinstance(node.func, ast.Name):
        raise TypeError("Please provide a valid form for prior: 'distribution(*args, **kwargs)'. "
                        "Provided: '{}'".format(expression))
    args = [ast.literal_eval(a) for a in node.args]
    kwargs = {k.arg: ast.literal_eval(k.value) for k in node.keywords}
    return node.func.id, args, kwargs


def build_dimension(name, expression):
    prior, args, kwargs = _literal_call(expression)
    klass = Integer if kwargs.pop('discrete', False) else Real
    if prior in ('choices', 'enum'):
        if not args:
            raise TypeError("Parameter '{}': Expected argument with categories.".format(name))
        cats = args[0] if isinstance(args[0], (dict, list)) else args
        return Categorical(name, cats, **kwargs)
    if prior in ('uniform', 'random'):
        if len(args) == 2:
            return klass(name, 'uniform', args[0], args[1] - args[0], **kwargs)
        return klass(name, 'uniform', *args, **kwargs)
    if prior in ('normal', 'gaussian'):
        return klass(name, 'norm', *args, **kwargs)
    if prior == 'loguniform':
        return klass(name, 'reciprocal', *args, **kwargs)
    return klass(name, prior, *args, **kwargs)


def build_space(definitions):
    """``[('x', 'uniform(-5, 10)'), ...]`` or ``['-x~uniform(-5, 10)', ...]`` -> Space."""
    space = Space()
    for item in definitions:
        if isinstance(item, str):
            name, expr = item.lstrip('-').split('~', 1)
        else:
            name, expr = item
        space.register(build_dimension(name, expr))
    return space


def rosen2():
    return build_space(['-x~uniform(-5, 10)', '-y~uniform(-5, 10)'])


def real16():
    return build_space(['-u%d~uniform(-1, 1)' % i for i in range(8)] +
                       ['-l%d~loguniform(1e-5, 1)' % i for i in range(8)])


def mixed64():
    """16 uniform + 8 loguniform + 8 normal Real, 16 uniform Integer,
    16 Categorical x 8 -- 64 slots, 448 bytes per candidate."""
    return build_space(['-u%d~uniform(-1, 1)' % i for i in range(16)] +
                       ['-l%d~loguniform(1e-5, 1)' % i for i in range(8)] +
                       ['-g%d~normal(0, 1)' % i for i in range(8)] +
                       ['-i%d~uniform(-100, 100, discrete=True)' % i for i in range(16)] +
                       ['-c%d~choices([0, 1, 2, 3, 4, 5, 6, 7])' % i for i in range(16)])


def cat32x100():
    return build_space(['-c%d~choices(%r)' % (i, list(range(100))) for i in range(32)])


NAMED = {'rosen2': rosen2, 'real16': real16, 'mixed64': mixed64, 'cat32x100': cat32x100}
