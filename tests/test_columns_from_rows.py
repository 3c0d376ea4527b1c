"""Host-side conversion of points (lists of tuples) into columns -- ``Space.columns_from_rows`` -- which every
membership / transform / observe call of the compatibility interface goes through.  No device needed: the fast
path (scalar dimensions, full rows: a C-level transpose, one dtype probe per numeric column, a hashed category
look-up) must agree with the row-by-row loop it stands in for, and leave to that loop whatever is unusual
(strings offered to numeric dimensions raise the reference's TypeError, space.py:164; mis-shaped values make the
row False, space.py:162,420)."""
from collections import OrderedDict

import numpy as np
import pytest

from orion_b200.space import Categorical, Integer, Real, Space, _column_values, _index_of


def scalar_space():
    space = Space()
    space.register(Real('x', 'uniform', -1, 2))
    space.register(Integer('k', 'uniform', -5, 10))
    space.register(Categorical('c', ('a', 2, 2.5, None, 'b')))
    space.register(Categorical('p', OrderedDict((('u', 0.3), ('v', 0.7)))))
    return space


def slow_columns(space, points):
    """The row-by-row statement: ``_column_values`` per value."""
    dims = list(space.values())
    cols = [[] for _ in dims]
    for p in points:
        for j, dim in enumerate(dims):
            cols[j].append(_column_values(dim, p[j])[0])
    return cols


def test_fast_path_equals_the_row_by_row_loop():
    space = scalar_space()
    rng = np.random.RandomState(0)
    cats, cats2 = space['c'].categories, space['p'].categories
    points = []
    for i in range(500):
        x = [rng.uniform(-2, 2), np.float64(rng.uniform(-2, 2)), np.float32(0.5), 1, True][i % 5]
        k = [int(rng.randint(-7, 7)), np.int64(3), 2.0, 2.5, np.int32(-5)][i % 5]
        c = [cats[i % 5], 'zzz', 2.0, '2', 3][i % 5]           # 2.0 == 2 (index 1); '2' and 3 are not categories
        points.append((x, k, c, cats2[i % 2]))
    dims = list(space.values())
    fast = Space._columns_of_scalar_rows(dims, points, len(points))
    assert fast is not None
    slow = slow_columns(space, points)
    for f, s_ in zip(fast, slow):
        assert np.array_equal(np.asarray(f, dtype=np.float64), np.asarray(s_, dtype=np.float64))
    assert fast[0].dtype == np.float64 and fast[1].dtype == np.float64 and fast[2].dtype == np.int32
    cols, row_ok = space.columns_from_rows(points)
    assert row_ok.all() and all(np.array_equal(a, b) for a, b in zip(cols, fast))
    assert fast[2][1] == -1 and fast[2][2] == 1 and fast[2][3] == -1 and fast[2][4] == -1


def test_category_lookup_is_first_match_by_equality():
    dim = Categorical('c', (1, 1.0, True, '1', 0))              # 1 == 1.0 == True: all find index 0
    space = Space()
    space.register(dim)
    for value, want in ((1, 0), (1.0, 0), (True, 0), ('1', 3), (0, 4), (False, 4), (2, -1)):
        assert _index_of(dim, value) == want
        fast = Space._columns_of_scalar_rows([dim], [(value,)], 1)
        assert fast is not None and int(fast[0][0]) == want


def test_unusual_inputs_are_left_to_the_general_loop():
    space = scalar_space()
    dims = list(space.values())
    good = (0.5, 1, 'a', 'u')
    # a string offered to a numeric dimension: the general loop raises the reference's TypeError
    assert Space._columns_of_scalar_rows(dims, [good, ('0.5', 1, 'a', 'u')], 2) is None
    with pytest.raises(TypeError):
        space.columns_from_rows([good, ('0.5', 1, 'a', 'u')])
    # a mis-shaped value: row False, the others stand
    assert Space._columns_of_scalar_rows(dims, [good, ([0.5, 0.6], 1, 'a', 'u')], 2) is None
    cols, row_ok = space.columns_from_rows([good, ([0.5, 0.6], 1, 'a', 'u')])
    assert list(row_ok) == [True, False]
    # None where a number belongs (object dtype): not silently nan
    assert Space._columns_of_scalar_rows(dims, [good, (None, 1, 'a', 'u')], 2) is None
    # an unhashable value offered to a categorical
    assert Space._columns_of_scalar_rows(dims, [good, (0.5, 1, ['a'], 'u')], 2) is None
    cols, row_ok = space.columns_from_rows([good, (0.5, 1, ['a'], 'u')])
    assert list(row_ok) == [True, False]
    # wrong arity: not the fast path's business
    assert Space._columns_of_scalar_rows(dims, [good, (0.5, 1, 'a')], 2) is None


def test_unhashable_categories_fall_back():
    dim = Categorical('c', ((1, 2), [3], 'x'))
    space = Space()
    space.register(dim)
    assert Space._columns_of_scalar_rows([dim], [('x',)], 1) is None
    assert _index_of(dim, 'x') == 2 and _index_of(dim, [3]) == 1


def test_empty_list_of_points():
    space = scalar_space()
    cols, row_ok = space.columns_from_rows([])
    assert len(row_ok) == 0 and len(cols) == 4 and all(len(c) == 0 for c in cols)
