"""ob2_set_dedup_static: dedup for callers that cannot allocate on the device.

Some users of libob2 hold nothing but the input and output buffers they were
given -- a plugin inside a process whose framework owns all of HBM through its
own pool, a producer loop that must never hit cudaMalloc's device-wide
synchronisation, a minimal FFI binding.  For them the per-entry scratch of the
two-pass dedup is a fixed array inside the library instead of a caller buffer.

This file is collected last on purpose (``zz``): if it fails, a run that stops
at the first failure has already run everything else.  It calls the entry
point through the C ABI and requires bit-identical verdicts to the existing
``ob2_set_dedup`` (the bar every dedup test here uses), on two calls in a row:
the second call finds the first call's leftovers in the shared array.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _fingerprints(rng, n):
    """n random non-empty fingerprints as an int64 [n][2] array (lo != 0)."""
    fp = rng.randint(1, 2 ** 62, size=(n, 2), dtype=np.int64)
    return fp


def _twin_histories(n_slots, history_fp):
    """Two hash sets with the same contents: one per entry point."""
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    space = spaces.rosen2()
    sets = [TrialHistory(space, n_slots=n_slots) for _ in range(2)]
    dev_fp = torch.from_numpy(history_fp).cuda()
    for s in sets:
        s.insert_fingerprints(dev_fp)
    return sets


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dedup_caller_scratch(lib, hist, fp, order, n):
    from orion_b200 import _lib
    keep = torch.zeros(fp.shape[0], dtype=torch.uint8, device='cuda')
    scratch = torch.empty(max(n, 1), dtype=torch.int32, device='cuda')
    _lib.check(lib.ob2_set_dedup(hist.handle, n, fp.data_ptr(), order.data_ptr() if order is not None else None,
                                 scratch.data_ptr(), keep.data_ptr(), _stream()))
    return keep.cpu().numpy()


def _dedup_static(lib, hist, fp, order, n):
    from orion_b200 import _lib
    keep = torch.zeros(fp.shape[0], dtype=torch.uint8, device='cuda')      # the caller's output buffer
    _lib.check(lib.ob2_set_dedup_static(hist.handle, n, fp.data_ptr(),
                                        order.data_ptr() if order is not None else None,
                                        keep.data_ptr(), _stream()))
    return keep.cpu().numpy()


def test_static_scratch_dedup_matches_caller_scratch_dedup_twice():
    from orion_b200 import _lib
    lib = _lib.load()
    rng = np.random.RandomState(2024)
    history_fp = _fingerprints(rng, 20000)
    with_scratch, static = _twin_histories(1 << 18, history_fp)

    # call 1: fresh rows, copies of history rows, copies of earlier rows of the batch
    n1 = 50000
    fp1 = _fingerprints(rng, n1)
    take = rng.random_sample(n1) < 0.15
    fp1[take] = history_fp[rng.randint(0, len(history_fp), size=int(take.sum()))]
    for i in np.nonzero(rng.random_sample(n1) < 0.10)[0]:
        if i:
            fp1[i] = fp1[rng.randint(0, i)]
    d1 = torch.from_numpy(fp1).cuda()
    want1 = _dedup_caller_scratch(lib, with_scratch, d1, None, n1)
    got1 = _dedup_static(lib, static, d1, None, n1)
    assert np.array_equal(got1, want1)
    assert 0 < int(want1.sum()) < n1
    assert len(static) == len(with_scratch)

    # call 2, straight after: shorter than call 1 (the array still holds call 1's
    # tail), explicit order keys, padding entries, copies of rows call 1 kept
    n2 = 30000
    fp2 = _fingerprints(rng, n2)
    again = rng.random_sample(n2) < 0.20
    fp2[again] = fp1[rng.randint(0, n1, size=int(again.sum()))]
    fp2[rng.random_sample(n2) < 0.05] = 0                                   # padding: lo == 0
    for i in np.nonzero(rng.random_sample(n2) < 0.10)[0]:
        if i:
            fp2[i] = fp2[rng.randint(0, i)]
    order2 = torch.from_numpy(rng.permutation(n2).astype(np.int32)).cuda()  # first-wins by key, not position
    d2 = torch.from_numpy(fp2).cuda()
    want2 = _dedup_caller_scratch(lib, with_scratch, d2, order2, n2)
    got2 = _dedup_static(lib, static, d2, order2, n2)
    assert np.array_equal(got2, want2)
    assert 0 < int(want2.sum()) < n2
    assert len(static) == len(with_scratch)

    # and both sets ended up holding the same fingerprints
    probe = torch.from_numpy(np.concatenate([fp1, fp2[fp2[:, 0] != 0]])).cuda()
    assert bool(static.probe_fingerprints(probe).all().item())
    assert bool(with_scratch.probe_fingerprints(probe).all().item())


def test_static_scratch_dedup_leaves_free_device_memory_untouched():
    """The point of the entry point: between the moment the set exists and the
    moment the verdicts are there, the amount of free device memory (the
    driver's cudaMemGetInfo figure, not torch's pool statistics) must not move
    -- not even on the very first call in a process, because creating the set
    is what loads the library's hash-set module and finds its built-in scratch.
    Run alone in a fresh process this IS the first call; after the other tests
    it is a later one; both must hold."""
    from orion_b200 import _lib
    lib = _lib.load()
    rng = np.random.RandomState(7)
    hist = _twin_histories(1 << 18, _fingerprints(rng, 1000))[1]       # ob2_set_create: module load + look-up
    n = 20000
    fps = [torch.from_numpy(_fingerprints(rng, n)).cuda() for _ in range(3)]
    keep = torch.zeros(n, dtype=torch.uint8, device='cuda')            # every buffer exists before measuring
    stream = _stream()
    torch.cuda.synchronize()
    free_before, _total = torch.cuda.mem_get_info()
    for fp in fps:
        _lib.check(lib.ob2_set_dedup_static(hist.handle, n, fp.data_ptr(), None, keep.data_ptr(), stream))
        torch.cuda.current_stream().synchronize()
        free_now, _total = torch.cuda.mem_get_info()
        assert free_now == free_before, "free device memory moved by %d bytes across a static dedup call" % (
            free_before - free_now)
    assert int(keep.sum().item()) == n                                  # fresh rows: all kept


def test_static_scratch_dedup_rejects_oversized_batches():
    from orion_b200 import _lib
    lib = _lib.load()
    hist = _twin_histories(1 << 10, _fingerprints(np.random.RandomState(1), 8))[0]
    n = _lib.STATIC_DEDUP_MAX + 1
    fp = torch.ones((n, 2), dtype=torch.int64, device='cuda')
    keep = torch.zeros(n, dtype=torch.uint8, device='cuda')
    rc = lib.ob2_set_dedup_static(hist.handle, n, fp.data_ptr(), None, keep.data_ptr(), _stream())
    assert rc == -3                                                          # OB2_ETOOBIG
    msg = lib.ob2_last_error().decode()
    assert 'exceeds the built-in scratch' in msg and str(_lib.STATIC_DEDUP_MAX) in msg
    assert not bool(keep.any().item())                                       # nothing ran
    # the Python wrapper surfaces the same error
    with pytest.raises(_lib.Ob2Error):
        hist.dedup_fingerprints_static(fp, keep)
    # at the limit it is accepted
    hist_big = _twin_histories(1 << 22, _fingerprints(np.random.RandomState(2), 8))[0]
    fp_ok = torch.from_numpy(_fingerprints(np.random.RandomState(3), _lib.STATIC_DEDUP_MAX)).cuda()
    keep_ok = torch.zeros(_lib.STATIC_DEDUP_MAX, dtype=torch.uint8, device='cuda')
    hist_big.dedup_fingerprints_static(fp_ok, keep_ok)
    assert int(keep_ok.sum().item()) == len(hist_big) - 8
