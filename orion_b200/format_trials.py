"""Points <-> trial records, batch-first.

Mirror of ``src/orion/core/utils/format_trials.py:16-55`` (``trial_to_tuple``,
``tuple_to_trial``) for the step right after the hot path:
``Producer.produce`` maps every suggested point to a ``Trial`` and registers it
(``src/orion/core/worker/producer.py:47-51``), one Python object per parameter.
Here the deduplicated survivors are gathered on the device (``ob2_compact``),
copied to the host once, and expanded into the same ``params`` records --
``{'name', 'type', 'value'}`` per dimension, the shape ``Trial.to_dict()``
stores (``src/orion/core/worker/trial.py:136-141, 170-191``).

The reference's ``Trial`` class belongs to its storage layer (out of scope
here); a trial is represented by the plain dict that class serialises to.
"""
import numpy


def tuple_to_trial(data, space):
    """One point (tuple in space order) -> trial record with only ``params`` filled."""
    assert len(data) == len(space)
    dims = list(space.values())
    return {'params': [dict(name=dims[j].name, type=dims[j].type, value=data[j]) for j in range(len(dims))]}


def trial_to_tuple(trial, space):
    """Trial record -> point; parameter order must be the space's."""
    params = trial['params']
    assert len(params) == len(space)
    for j, param in enumerate(params):
        assert space[j].name == param['name']
    return tuple(param['value'] for param in params)


def batch_to_trials(space, batch, keep=None):
    """Surviving rows of a device ColumnBatch -> list of trial records.

    ``keep`` defaults to ``batch.keep`` when the batch went through dedup, and
    to every row otherwise.  Survivors are compacted on the device and cross
    PCIe once.
    """
    keep = batch.keep if keep is None else keep
    # (wire: Integer / Categorical columns cross PCIe at their narrowest lossless width)
    dense = batch.compact(keep, wire=not batch.transformed) if keep is not None else batch
    rows = space.rows_from_columns(dense.to_numpy(), dense.n)
    return [tuple_to_trial(row, space) for row in rows]


def trials_to_batch(space, trials):
    """Completed trials -> device ColumnBatch (history ingestion:
    ``Producer.update`` -> ``trial_to_tuple`` -> ``observe``,
    ``src/orion/core/worker/producer.py:54-67``)."""
    from .columns import ColumnBatch
    dims = list(space.values())
    names = [d.name for d in dims]
    points = []
    for trial in trials:                        # trial_to_tuple for every trial, the space looked up once
        params = trial['params']
        assert len(params) == len(dims) and [p['name'] for p in params] == names
        points.append(tuple(p['value'] for p in params))
    cols, row_ok = space.columns_from_rows(points)
    if not row_ok.all():
        raise ValueError("a trial does not match the shape of the space")
    compiled = space.compiled()
    typed = [numpy.asarray(c, dtype={'f8': numpy.float64, 'i8': numpy.int64, 'i4': numpy.int32}[dt])
             for c, (_u, dt) in zip(cols, compiled.col_meta)]
    return ColumnBatch.from_numpy(compiled, typed)


def columns_to_record_batch(space, host_cols):
    """Host columns (one per scalar slot, space order; narrow wire dtypes
    welcome) -> ``pyarrow.RecordBatch``: the columnar form of
    ``[tuple_to_trial(p) for p in points]`` without one Python object per
    parameter.  Real -> float64, Integer -> the column's integer type,
    Categorical -> dictionary-encoded categories; a dimension with ``shape``
    contributes one field per element, named ``name[i]`` in C order.  The
    dimension types travel in the schema metadata (``b'orion.types'``)."""
    import json
    import pyarrow as pa
    arrays, names, types = [], [], {}
    s = 0
    for dim in space.values():
        k = int(numpy.prod(dim.shape, dtype=numpy.int64)) if dim.shape else 1
        types[dim.name] = dim.type
        for e in range(k):
            col = numpy.ascontiguousarray(host_cols[s + e])
            if dim.type == 'categorical':
                cats = list(dim.categories)
                try:
                    dictionary = pa.array(cats)
                except (pa.ArrowInvalid, pa.ArrowTypeError):      # categories of mixed Python types
                    dictionary = pa.array([repr(c) for c in cats])
                arr = pa.DictionaryArray.from_arrays(pa.array(col.astype(numpy.int32, copy=False)), dictionary)
            else:
                arr = pa.array(col)                                # zero-copy for primitive columns
            arrays.append(arr)
            names.append(dim.name if not dim.shape else '%s[%d]' % (dim.name, e))
        s += k
    schema = pa.schema([pa.field(n, a.type) for n, a in zip(names, arrays)],
                       metadata={b'orion.types': json.dumps(types).encode()})
    return pa.RecordBatch.from_arrays(arrays, schema=schema)


def batch_to_record_batch(space, batch, keep=None):
    """Surviving rows of a device ColumnBatch as one ``pyarrow.RecordBatch``:
    survivors are gathered on the device in the narrowed wire layout
    (``ob2_compact_wire``), copied to the host once, and wrapped column by
    column -- no per-row Python work (the bulk counterpart of
    :func:`batch_to_trials`)."""
    keep = batch.keep if keep is None else keep
    dense = batch.compact(keep, wire=not batch.transformed) if keep is not None else batch
    return columns_to_record_batch(space, dense.to_numpy(natural=False))


def record_batch_to_columns(space, record_batch):
    """Inverse of :func:`columns_to_record_batch`: host columns in the
    natural dtypes (float64 / int64 / int32 category index), ready for
    ``ColumnBatch.from_numpy`` -- bulk history ingestion (resume) without
    going through one trial record per row."""
    import pyarrow as pa
    cols = []
    f = 0
    for dim in space.values():
        k = int(numpy.prod(dim.shape, dtype=numpy.int64)) if dim.shape else 1
        for e in range(k):
            field = record_batch.schema.field(f)
            want = dim.name if not dim.shape else '%s[%d]' % (dim.name, e)
            if field.name != want:
                raise ValueError("field %d is %r, the space expects %r" % (f, field.name, want))
            arr = record_batch.column(f)
            if dim.type == 'categorical':
                if not pa.types.is_dictionary(arr.type):
                    raise ValueError("categorical dimension %r needs a dictionary-encoded field" % dim.name)
                cats = list(dim.categories)
                try:
                    expected = pa.array(cats).to_pylist()
                except (pa.ArrowInvalid, pa.ArrowTypeError):
                    expected = [repr(c) for c in cats]
                if arr.dictionary.to_pylist() != expected:
                    raise ValueError("dictionary of %r does not match the dimension's categories" % field.name)
                cols.append(arr.indices.to_numpy(zero_copy_only=False).astype(numpy.int32, copy=False))
            elif dim.type == 'integer':
                cols.append(arr.to_numpy(zero_copy_only=False).astype(numpy.int64, copy=False))
            else:
                cols.append(arr.to_numpy(zero_copy_only=False).astype(numpy.float64, copy=False))    # (a view of the arrow buffer)
            f += 1
    if f != record_batch.num_columns:
        raise ValueError("record batch has %d fields, the space %d slots" % (record_batch.num_columns, f))
    return cols
