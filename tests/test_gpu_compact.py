"""ob2_compact (survivor gather) and the trial-record encoder built on it
(src/orion/core/utils/format_trials.py:16-42 applied to whole batches)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('n', [1, 2, 511, 512, 513, 100_003])
@pytest.mark.parametrize('transformed', [False, True])
def test_compact_matches_boolean_indexing(n, transformed):
    from orion_b200 import spaces
    from orion_b200.transformer import build_required_space
    space = spaces.mixed64()
    compiled = build_required_space('real', space).compiled()
    tb = compiled.empty_columns(n, transformed=True)
    batch = compiled.sample(n, 9, transformed_out=tb)
    src = tb if transformed else batch
    rng = np.random.RandomState(n)
    for frac in (0.0, 0.5, 0.89, 1.0):
        keep_host = (rng.random_sample(n) < frac).astype(np.uint8)
        keep = torch.from_numpy(keep_host).cuda()
        dense = src.compact(keep)
        assert dense.n == int(keep_host.sum())
        for got, want in zip(dense.to_numpy(), src.to_numpy()):
            assert got.dtype == want.dtype and np.array_equal(got, want[keep_host.astype(bool)])


def test_batch_to_trials_round_trip():
    from collections import OrderedDict
    from orion_b200 import format_trials
    from orion_b200.history import TrialHistory
    from orion_b200.space import Categorical, Integer, Real, Space
    space = Space()
    space.register(Categorical('yolo', OrderedDict(zip(('asdfa', 2, 3, 4), (0.1, 0.2, 0.3, 0.4))), shape=2))
    space.register(Integer('yolo2', 'uniform', -3, 6))
    space.register(Real('yolo3', 'norm', 0.9))
    # tuple_to_trial / trial_to_tuple: tests/unittests/core/test_utils_format.py
    point = (np.array(['asdfa', 2], dtype=object), 0, 3.5)
    trial = format_trials.tuple_to_trial(point, space)
    assert [p['name'] for p in trial['params']] == ['yolo', 'yolo2', 'yolo3']
    assert [p['type'] for p in trial['params']] == ['categorical', 'integer', 'real']
    assert format_trials.trial_to_tuple(trial, space)[1:] == (0, 3.5)
    with pytest.raises(AssertionError):
        format_trials.tuple_to_trial(point[:2], space)
    # batch path: dedup -> compact -> one host copy -> records
    hist = TrialHistory(space, capacity=4096)
    counters = torch.from_numpy(np.random.RandomState(0).randint(0, 300, size=500).astype(np.int64)).cuda()
    batch = space.compiled().sample(500, 3, index=counters, with_fingerprint=True)    # repeated counters = repeated rows
    hist.dedup(batch)
    trials = format_trials.batch_to_trials(space, batch)
    kept = int(batch.keep[:500].sum().item())
    assert len(trials) == kept == len(np.unique(counters.cpu().numpy()))
    rows = space.rows_from_columns(batch.to_numpy(), 500)
    keep = batch.keep[:500].cpu().numpy().astype(bool)
    want = [r for r, k in zip(rows, keep) if k]
    for t, w in zip(trials, want):
        got = format_trials.trial_to_tuple(t, space)
        assert np.all(got[0] == w[0]) and got[1:] == w[1:]
    # history ingestion of stored trials: they are all known now
    again = format_trials.trials_to_batch(space, trials)
    assert not bool(hist.dedup(again).any().item())


def _width_space():
    from orion_b200.space import Categorical, Integer, Real, Space
    space = Space()
    space.register(Integer('i1', 'uniform', -128, 256))            # [-128, 128)   -> int8
    space.register(Integer('i2a', 'uniform', -128, 257))           # [-128, 129)   -> int16
    space.register(Integer('i2b', 'reciprocal', 1, 1000))          # [1, 1000)     -> int16
    space.register(Real('x', 'uniform', -1, 2))
    space.register(Integer('i4', 'uniform', 0, 1000000))           # -> int32
    space.register(Integer('i8', 'norm', 0, 1e12))                 # unbounded     -> int64
    space.register(Categorical('c1', list(range(256))))            # -> uint8
    space.register(Categorical('c2', list(range(257)), shape=2))   # -> uint16
    space.register(Integer('neg', 'uniform', -40000, 10))          # [-40000, -39990) -> int32
    return space


def test_wire_layout_is_the_narrowest_lossless_one():
    compiled = _width_space().compiled()
    assert [dt for _u, dt in compiled.wire_meta] == ['i1', 'i2', 'i2', 'f8', 'i4', 'i8', 'u1', 'u2', 'u2', 'i4']
    assert compiled.wire_row_bytes == 1 + 2 + 2 + 8 + 4 + 8 + 1 + 2 + 2 + 4
    units = [u for u, _dt in compiled.wire_meta]
    assert units == [0, 1, 3, 5, 13, 17, 25, 26, 28, 30]
    from orion_b200 import spaces
    assert spaces.mixed64().compiled().wire_row_bytes == 32 * 8 + 16 + 16       # 288 instead of 448


@pytest.mark.parametrize('n', [1, 2, 513, 100_003])
def test_compact_wire_matches_natural_compaction(n):
    from orion_b200 import spaces
    for space in (_width_space(), spaces.mixed64()):
        compiled = space.compiled()
        batch = compiled.sample(n, 17)
        assert bool(batch.mask[:n].all())
        rng = np.random.RandomState(n)
        for frac in (0.0, 0.89, 1.0):
            keep_host = (rng.random_sample(n) < frac).astype(np.uint8)
            keep = torch.from_numpy(keep_host).cuda()
            wide = batch.compact(keep)
            wire = batch.compact(keep, wire=True)
            assert wire.wire and wire.n == wide.n == int(keep_host.sum())
            assert wire.buf.numel() == compiled.wire_row_bytes * wire.n_stride
            raw = wire.to_numpy(natural=False)
            for (u, dt), col in zip(compiled.wire_meta, raw):
                assert col.dtype == np.dtype(dt)
            for got, want, narrow in zip(wire.to_numpy(), wide.to_numpy(), raw):
                assert got.dtype == want.dtype and np.array_equal(got, want)
                assert np.array_equal(narrow.astype(want.dtype), want)            # lossless
            for j in range(len(raw)):                                             # typed device views too
                assert np.array_equal(wire.column(j).cpu().numpy(), raw[j])


def test_batch_to_record_batch_equals_trial_records():
    """The columnar hand-over (pyarrow) carries exactly the rows batch_to_trials builds."""
    from orion_b200 import format_trials, spaces
    from orion_b200.history import TrialHistory
    space = spaces.mixed64()
    n = 2000
    hist = TrialHistory(space, capacity=8 * n)
    counters = torch.from_numpy(np.random.RandomState(1).randint(0, n // 2, size=n).astype(np.int64)).cuda()
    batch = space.compiled().sample(n, 3, index=counters, with_fingerprint=True)
    hist.dedup(batch)
    rb = format_trials.batch_to_record_batch(space, batch)
    trials = format_trials.batch_to_trials(space, batch)
    assert rb.num_rows == len(trials) == int(batch.keep[:n].sum().item())
    assert rb.schema.names == list(space.keys())
    table = rb.to_pydict()
    for j, name in enumerate(space.keys()):
        want = [t['params'][j]['value'] for t in trials]
        got = table[name]
        if space[name].type == 'real':
            assert np.array_equal(np.asarray(got), np.asarray(want, dtype=np.float64))
        else:
            assert [int(g) for g in got] == [int(w) for w in want]
    # narrow wire integers arrive as narrow arrow types: nothing was widened on the way
    import pyarrow as pa
    assert rb.schema.field('i0').type == pa.int8() and rb.schema.field('u0').type == pa.float64()
    # ... and back: the record batch re-enters a fresh history as exactly those rows (bulk resume)
    from orion_b200.columns import ColumnBatch
    again = ColumnBatch.from_numpy(space.compiled(), format_trials.record_batch_to_columns(space, rb))
    fresh = TrialHistory(space, capacity=8 * n)
    fresh.observe(again)
    assert len(fresh) == rb.num_rows == len(hist)
    assert int(fresh.dedup(space.compiled().sample(n, 3, index=counters, with_fingerprint=True)).sum().item()) == 0


@pytest.mark.parametrize('wire', [False, True])
def test_columns_to_host_copies_the_used_rows_of_every_column(wire):
    """ob2_columns_to_host: the survivors' rows into a pinned host block of the same layout with
    a few 2-D copies; equal to what to_numpy reads back, for the original, the wire and the
    transformed layout, including a host stride different from the device stride."""
    import torch
    from orion_b200 import spaces
    from orion_b200.columns import ColumnBatch, _NUMPY, _SIZE
    from orion_b200.transformer import build_required_space
    space = spaces.mixed64()
    compiled = space.compiled()
    n = 5000
    batch = compiled.sample(n, 11, with_mask=True)
    keep = (torch.arange(batch.n_stride, device='cuda') % 3 != 0).to(torch.uint8)
    dense = batch.compact(keep=keep, wire=wire)
    k = dense.n
    assert 0 < k < n
    want = dense.to_numpy(natural=False)
    for host_stride in (dense.n_stride, 4096):
        row = compiled.wire_row_bytes if wire else compiled.row_bytes
        host = torch.full((row * host_stride,), 255, dtype=torch.uint8).pin_memory()
        dense.copy_rows_to_host(host, k, host_stride=host_stride)
        torch.cuda.synchronize()
        h = host.numpy()
        for (units, dt), w in zip(dense._meta, want):
            got = h[units * host_stride:units * host_stride + _SIZE[dt] * k].view(_NUMPY[dt])
            assert np.array_equal(got, w)
        # nothing beyond the k rows of a column was touched
        units, dt = dense._meta[0]
        assert (h[units * host_stride + _SIZE[dt] * k:units * host_stride + _SIZE[dt] * host_stride] == 255).all()
    if not wire:        # transformed layout (one-hot columns are a run of 8-byte columns)
        tc = build_required_space('real', space).compiled()
        _b, tb, _k = tc.suggest(1000, 3)
        host = torch.zeros(tc.trow_bytes * tb.n_stride, dtype=torch.uint8).pin_memory()
        tb.copy_rows_to_host(host, 1000)
        torch.cuda.synchronize()
        assert torch.equal(host[:tb.buf.numel()].reshape(-1, tb.n_stride * 8)[:, :8000],
                           tb.buf.cpu().reshape(-1, tb.n_stride * 8)[:, :8000])
