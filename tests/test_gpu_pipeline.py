"""The stream pipeline gives exactly the results of the one-stream path."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_pipeline_matches_serial_path():
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    from orion_b200.pipeline import CandidatePipeline
    from orion_b200.transformer import build_required_space
    space = spaces.mixed64()
    compiled = build_required_space('real', space).compiled()
    n, steps, seed = 100_000, 6, 42
    rng = np.random.RandomState(0)
    # counters with repeats inside a batch and across batches
    index = [torch.from_numpy(rng.randint(0, 3 * n, size=n).astype(np.int64)).cuda() for _ in range(steps)]

    serial_hist = TrialHistory(compiled, capacity=2 * n * steps)
    want = []
    for k in range(steps):
        tb = compiled.empty_columns(n, transformed=True)
        b = compiled.sample(n, seed, index=index[k], with_fingerprint=True, transformed_out=tb)
        back = compiled.reverse(tb)
        keep = serial_hist.dedup(b).clone()
        want.append((b.buf.clone(), tb.buf.clone(), back.buf.clone(), keep, b.mask[:n].clone()))

    pipe_hist = TrialHistory(compiled, capacity=2 * n * steps)
    pipe = CandidatePipeline(compiled, pipe_hist, n, depth=2)
    got = []
    for k in range(steps):
        slot = pipe.submit(seed, index=index[k])
        pipe.drain()                # results of this batch are now ordered before the copies below
        got.append((slot.batch.buf.clone(), slot.tbatch.buf.clone(), slot.back.buf.clone(),
                    slot.batch.keep[:n].clone(), slot.batch.mask[:n].clone()))
    torch.cuda.synchronize()
    for k, (w, g) in enumerate(zip(want, got)):
        for a, b in zip(w, g):
            assert torch.equal(a, b), 'step %d' % k
    assert len(pipe_hist) == len(serial_hist)

    # without draining between submissions (real overlap) the verdicts are the same
    free_hist = TrialHistory(compiled, capacity=2 * n * steps)
    free = CandidatePipeline(compiled, free_hist, n, depth=2)
    kept = []
    for k in range(steps):
        slot = free.submit(seed, index=index[k])
        with torch.cuda.stream(free.s_dedup):       # ordered after this batch's dedup, before the slot is reused
            kept.append(slot.batch.keep[:n].sum())
    free.drain()
    torch.cuda.synchronize()
    assert [int(x.item()) for x in kept] == [int(w[3].sum().item()) for w in want]
    assert len(free_hist) == len(serial_hist)
