"""The prior codes the host writes into slot descriptors (orion_b200/descriptor.py) are the enumerators of
include/ob2.h: every supported scipy name must compile to the code the kernels switch on."""
import os
import re

import pytest

from orion_b200 import descriptor
from orion_b200.space import Integer, Real

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ALIASES = {'loguniform': 'RECIPROCAL', 'erlang': 'GAMMA', 'fisk': 'BURR', 'maxwell': 'CHI', 'semicircular': 'RDIST',
           'bernoulli': 'BINOM'}

# scipy name -> arguments that pass scipy's own validation
EXAMPLES = {
    'uniform': (0, 1), 'reciprocal': (1e-3, 1.0), 'loguniform': (1e-3, 1.0), 'norm': (0, 1),
    'expon': (), 'lognorm': (0.5,), 'logistic': (), 'laplace': (), 'gumbel_r': (), 'weibull_min': (1.5,), 'pareto': (2.5,),
    'rayleigh': (), 'halfnorm': (), 'powerlaw': (3.0,), 'alpha': (0.9,), 'randint': (-3, 5), 'poisson': (3.0,),
    'gamma': (2.0,), 'erlang': (3,), 'beta': (2.0, 5.0), 'cauchy': (), 'chi2': (3.0,), 't': (2.5,), 'truncnorm': (-1.0, 2.0),
    'geom': (0.3,), 'binom': (20, 0.3),
    'gumbel_l': (), 'genpareto': (0.1,), 'lomax': (2.0,), 'burr': (2.0, 1.5), 'fisk': (3.0,), 'burr12': (2.0, 1.5),
    'invweibull': (2.5,), 'truncexpon': (4.0,), 'bradford': (2.0,), 'arcsine': (), 'triang': (0.3,), 'loglaplace': (2.0,),
    'exponpow': (1.5,), 'gompertz': (0.7,), 'genlogistic': (0.8,), 'hypsecant': (), 'halfcauchy': (), 'halflogistic': (),
    'chi': (3.5,), 'maxwell': (), 'invgamma': (3.0,), 'nakagami': (1.7,), 'gibrat': (), 'genextreme': (0.2,),
    'kappa3': (1.5,), 'dweibull': (1.3,), 'anglit': (), 'exponweib': (2.0, 1.5), 'mielke': (2.0, 3.0),
    'genhalflogistic': (0.7,), 'tukeylambda': (0.3,), 'levy': (), 'levy_l': (), 'weibull_max': (1.8,), 'rdist': (4.0,),
    'semicircular': (), 'dgamma': (1.5,), 'gennorm': (1.5,), 'moyal': (), 'wrapcauchy': (0.4,),
    'f': (5.0, 7.0), 'loggamma': (0.5,), 'nbinom': (5, 0.3), 'hypergeom': (50, 10, 20), 'vonmises': (2.0,),
    'johnsonsb': (0.5, 1.5), 'johnsonsu': (-0.5, 2.0), 'fatiguelife': (0.7,), 'powernorm': (2.5,), 'powerlognorm': (2.0, 0.5),
    'gengamma': (2.0, 1.5), 'kappa4': (0.5, 0.3), 'laplace_asymmetric': (1.7,), 'trapezoid': (0.2, 0.7),
    'truncpareto': (2.0, 5.0), 'truncweibull_min': (1.5, 0.5, 3.0), 'skewcauchy': (0.3,), 'bernoulli': (0.3,),
    'planck': (0.5,), 'dlaplace': (0.8,), 'boltzmann': (1.4, 19),
}
DISCRETE = ('randint', 'poisson', 'geom', 'binom', 'nbinom', 'hypergeom', 'bernoulli', 'planck', 'dlaplace', 'boltzmann')


def header_codes():
    text = open(os.path.join(ROOT, 'include', 'ob2.h')).read()
    return {name: int(value) for name, value in re.findall(r'OB2_DIST_([A-Z0-9_]+) = (\d+)', text)}


def test_every_supported_prior_compiles_to_the_header_s_code():
    codes = header_codes()
    assert codes['LAST'] == max(v for k, v in codes.items() if k != 'LAST')
    seen = set()
    for name, args in EXAMPLES.items():
        dim = (Integer if name in DISCRETE else Real)('x', name, *args)
        code = descriptor._prior_params(dim)[0]
        want = codes[ALIASES.get(name, name.upper())]
        assert code == want, (name, code, want)
        seen.add(code)
    # ... and every code of the header (CONST / NONE aside) is reachable from some scipy name
    assert seen == {v for k, v in codes.items() if k not in ('LAST', 'CONST', 'NONE')}


def test_python_constants_match_the_header():
    from orion_b200 import _lib
    codes = header_codes()
    assert (_lib.DIST_UNIFORM, _lib.DIST_RECIPROCAL, _lib.DIST_NORM, _lib.DIST_CONST, _lib.DIST_NONE) == \
        (codes['UNIFORM'], codes['RECIPROCAL'], codes['NORM'], codes['CONST'], codes['NONE'])
    from orion_b200 import mt_stream
    assert mt_stream._UNIFORM_FED == (codes['EXPON'], codes['LOGISTIC'], codes['LAPLACE'])


def test_scale_zero_and_unknown_priors():
    from orion_b200 import _lib
    assert descriptor._prior_params(Real('x', 'uniform', 3, 0))[0] == _lib.DIST_CONST
    assert descriptor._prior_params(Real('x', 'ncx2', 3.0, 1.0))[0] == _lib.DIST_NONE
    with pytest.raises(TypeError):
        descriptor._prior_params(Real('x', 'norm', 0, 1, 2, 3))
