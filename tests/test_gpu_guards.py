"""Out-of-bounds guard: every output buffer of the hot-path kernels is carved
out of a larger canary-filled allocation; after the call nothing outside the
buffer may have changed (sizes chosen to exercise odd tails and partial
blocks).  compute-sanitizer is not available on the GPU pool, so this is the
bounds check the kernels get."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 4096
CANARY = 0xAB


class Guarded(object):
    def __init__(self, nbytes, dev):
        self.nbytes = nbytes
        self.raw = torch.full((nbytes + 2 * GUARD,), CANARY, dtype=torch.uint8, device=dev)
        self.buf = self.raw[GUARD:GUARD + nbytes]
        assert self.buf.data_ptr() % 256 == 0

    def view(self, dtype, shape):
        return self.buf.view(dtype).view(shape)

    def intact(self):
        return bool((self.raw[:GUARD] == CANARY).all()) and bool((self.raw[GUARD + self.nbytes:] == CANARY).all())


def _guarded_batch(compiled, n, transformed=False, wire=False):
    from orion_b200.columns import ColumnBatch
    from orion_b200.descriptor import round_up
    ns = round_up(max(n, 1))
    row = compiled.wire_row_bytes if wire else compiled.trow_bytes if transformed else compiled.row_bytes
    g = Guarded(row * ns, compiled.device)
    return ColumnBatch(compiled, n, ns, g.buf, transformed, wire=wire), g


@pytest.mark.parametrize('n', [1, 2, 31, 33, 511, 513, 1025, 4097])
@pytest.mark.parametrize('requirement', ['real', 'integer'])
def test_no_kernel_writes_outside_its_buffers(n, requirement):
    from orion_b200.history import TrialHistory
    from orion_b200.transformer import build_required_space
    from tests import golden_util as gu
    g0 = gu.load('bounded_reject')
    extra = [dict(d, name='m_' + d['name']) for d in gu.load('mixed_scalar')['dims'][5:]]
    dims = g0['dims'] + extra + [dict(name='sh', kind='categorical', categories=[1, 2, 3], shape=[2])]
    space = gu.product_space(dims)
    compiled = build_required_space(requirement, space).compiled()
    dev = compiled.device
    guards = []

    def batch(transformed=False, wire=False):
        b, g = _guarded_batch(compiled, n, transformed, wire)
        guards.append(g)
        return b

    def side(b, fp=False, mask=False, keep=False):
        ns = b.n_stride
        if fp:
            g = Guarded(ns * 16, dev)
            guards.append(g)
            b.fp = g.view(torch.int64, (ns, 2))
        if mask:
            g = Guarded(ns, dev)
            guards.append(g)
            b.mask = g.buf
        if keep:
            g = Guarded(ns, dev)
            guards.append(g)
            b.keep = g.buf
        gf = Guarded(compiled.n_slots * 4, dev)
        guards.append(gf)
        b.n_failed = gf.view(torch.int32, (compiled.n_slots,))
        return b

    # sample / sample+transform / suggest
    plain = compiled.sample(n, 3, with_fingerprint=True, out=side(batch(), fp=True, mask=True))
    tb = batch(transformed=True)
    fused = compiled.sample(n, 3, with_fingerprint=True, out=side(batch(), fp=True, mask=True), transformed_out=tb)
    sb, st, sk = compiled.suggest(n, 3, out=side(batch(), mask=True), transformed_out=batch(transformed=True),
                                  back_out=side(batch(), fp=True, keep=True))
    # transform / reverse (+ fingerprint) / fingerprint / contains
    t2 = compiled.transform(plain, out=batch(transformed=True))
    r2 = compiled.reverse(t2, out=batch())
    r3 = compiled.reverse(t2, out=side(batch(), fp=True), with_fingerprint=True)
    compiled.fingerprint(side(r2, fp=True))
    compiled.contains(r2)
    # dedup + both compactions (scratch and count guarded too)
    hist = TrialHistory(compiled, capacity=4 * n + 64)
    hist.dedup(sk)
    gs = Guarded(max(int(compiled.lib.ob2_compact_scratch_bytes(n)), 8), dev)
    gc = Guarded(8, dev)
    guards.extend([gs, gc])
    for wire in (False, True):
        sk.compact(out=batch(wire=wire), count=gc.view(torch.int64, (1,)), scratch=gs.buf, wire=wire)
    st.compact(keep=sk.keep, out=batch(transformed=True), count=gc.view(torch.int64, (1,)), scratch=gs.buf)
    torch.cuda.synchronize()
    assert all(g.intact() for g in guards)
    # (and the guarded buffers hold the same results as ordinary ones)
    for a, b in zip(sb.to_numpy(), fused.to_numpy()):
        assert np.array_equal(a, b)
    assert torch.equal(sk.fp[:n], r3.fp[:n])


def test_misaligned_caller_pointers_are_refused_on_the_host():
    """A column block / fingerprint array offset by 8 bytes would fault inside the kernels
    (128-bit accesses) and poison the context; the entry points return OB2_EINVAL instead."""
    import ctypes
    import torch
    from orion_b200 import _lib, spaces
    lib = _lib.load()
    c = spaces.rosen2().compiled()
    n = 64
    buf = torch.zeros(c.row_bytes * n + 64, dtype=torch.uint8, device='cuda')
    fp = torch.zeros(2 * n + 8, dtype=torch.int64, device='cuda')
    mask = torch.zeros(n + 8, dtype=torch.uint8, device='cuda')
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    base = buf.data_ptr()
    assert lib.ob2_sample(c.handle, n, n, 1, 0, None, base + 8, None, None, None, st) == -1
    assert b'16-byte aligned' in lib.ob2_last_error() and b'cols' in lib.ob2_last_error()
    assert lib.ob2_sample(c.handle, n, n, 1, 0, None, base, None, fp.data_ptr() + 8, None, st) == -1
    assert lib.ob2_sample(c.handle, n, n, 1, 0, None, base, mask.data_ptr() + 1, None, None, st) == -1
    assert lib.ob2_fingerprint(c.handle, n, n, base, fp.data_ptr() + 8, st) == -1
    assert lib.ob2_contains(c.handle, n, n, base + 4, 0, mask.data_ptr(), st) == -1
    assert lib.ob2_transform(c.handle, n, n, base + 8, base, st) == -1
    # the aligned call still works afterwards: nothing reached the device
    assert lib.ob2_sample(c.handle, n, n, 1, 0, None, base, mask.data_ptr(), fp.data_ptr(), None, st) == 0
    torch.cuda.synchronize()
    assert bool(mask[:n].all().item())


def test_space_create_rejects_inconsistent_dimension_grouping():
    import ctypes
    from orion_b200 import _lib
    lib = _lib.load()

    def table(fix):
        arr = (_lib.SlotDesc * 2)()
        for s in range(2):
            arr[s].kind, arr[s].dist = _lib.KIND_REAL, _lib.DIST_UNIFORM
            arr[s].dim_first, arr[s].flags = 0, (_lib.SLOT_DIM_LAST if s == 1 else 0)
            arr[s].col_units = arr[s].t_units = 8 * s
            arr[s].t_op, arr[s].t_width = _lib.OP_IDENTITY, 1
            arr[s].scale, arr[s].high = 1.0, 1.0
        fix(arr)
        return arr

    out = ctypes.c_void_p()
    cdf = (ctypes.c_double * 1)(1.0)
    ok = table(lambda a: None)
    assert lib.ob2_space_create(ok, 2, cdf, 0, ctypes.byref(out)) == 0
    lib.ob2_space_destroy(out)

    def mixed_kinds(a):
        a[1].kind = _lib.KIND_INTEGER
    assert lib.ob2_space_create(table(mixed_kinds), 2, cdf, 0, ctypes.byref(out)) == -1
    assert b'inconsistent' in lib.ob2_last_error()

    def wrong_first(a):
        a[1].dim_first = 1
    assert lib.ob2_space_create(table(wrong_first), 2, cdf, 0, ctypes.byref(out)) == -1

    def never_closed(a):
        a[1].flags = 0
    assert lib.ob2_space_create(table(never_closed), 2, cdf, 0, ctypes.byref(out)) == -1
