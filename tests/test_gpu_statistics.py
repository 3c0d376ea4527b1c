"""Philox-mode marginals against the reference's distributions:
KS for continuous slots, chi-square for Integer / Categorical pmfs, N = 1e6 per
slot, alpha = 1e-3 with a Bonferroni correction over the slots; plus the
size-independent properties checked at BASELINE.json's full sizes."""
import numpy as np
import pytest
import torch
from scipy import stats

from tests.spec_util import specs_from_space

pytestmark = pytest.mark.gpu


def test_mixed64_marginals_match_reference_distributions():
    from orion_b200 import spaces
    space = spaces.mixed64()
    specs = specs_from_space(space)
    n = 1_000_000
    cols = space.compiled().sample(n, 0xC0FFEE).to_numpy()
    alpha = 1e-3 / len(cols)
    for s, (spec, x) in enumerate(zip(specs, cols)):
        if spec.kind == 'real':
            dist = getattr(stats, spec.prior)(*spec.args)
            p = stats.kstest(x, dist.cdf).pvalue
        elif spec.kind == 'integer':
            dist = getattr(stats, spec.prior)(*spec.args)
            lo, hi = int(x.min()), int(x.max())
            edges = np.arange(lo, hi + 2)
            expected = np.diff(dist.cdf(edges)) * n                 # pmf(k) = cdf(k+1) - cdf(k)
            observed = np.bincount(x - lo, minlength=hi - lo + 1)
            p = stats.chisquare(observed, expected * observed.sum() / expected.sum()).pvalue
            assert lo >= -100 and hi <= 99
        else:
            observed = np.bincount(x, minlength=len(spec.categories))
            p = stats.chisquare(observed, np.asarray(spec.probs) * n).pvalue
        assert p > alpha, 'slot %d (%s %s): p = %g' % (s, spec.kind, spec.prior, p)


def test_slots_and_candidates_are_uncorrelated():
    """Adjacent slots share a Philox block and adjacent candidates a thread:
    neither may correlate."""
    from orion_b200 import spaces
    cols = spaces.real16().compiled().sample(400_000, 7).to_numpy()
    u = np.stack(cols[:8])                                          # the 8 uniform(-1, 1) slots
    c = np.corrcoef(u)
    assert np.abs(c - np.eye(8)).max() < 0.01
    assert abs(np.corrcoef(u[0][:-1], u[0][1:])[0, 1]) < 0.01


def test_truncated_normal_keeps_its_shape_under_rejection():
    """Redraws must sample the prior conditioned on the bounds."""
    from orion_b200.space import Real, Space
    space = Space()
    space.register(Real('a', 'norm', 0, 1, low=-1.0, high=1.5))
    n = 500_000
    batch = space.compiled().sample(n, 5)
    ok = batch.mask[:n].cpu().numpy().astype(bool)
    x = batch.to_numpy()[0][ok]
    assert x.min() >= -1.0 and x.max() < 1.5
    p = stats.kstest(x, stats.truncnorm(-1.0, 1.5).cdf).pvalue
    assert p > 1e-4
    # 4 misses in a row (the reference's ValueError): P = (1 - 0.7745)^4 = 0.26 %
    missed = n - int(ok.sum())
    assert missed == int(batch.n_failed.sum().item())
    assert abs(missed / n - 0.2255 ** 4) < 5e-4


@pytest.mark.parametrize('name,n', [('real16', 1_000_000), ('mixed64', 10_000_000)])
def test_full_size_properties(name, n):
    """BASELINE.json configs 2 and 3 at full size: every candidate in space,
    transform('real') -> reverse is the identity, one-hot rows sum to one,
    fused == unfused, and re-running with the same key reproduces the batch."""
    from orion_b200 import spaces
    from orion_b200.transformer import build_required_space
    space = spaces.NAMED[name]()
    compiled = build_required_space('real', space).compiled()
    tb = compiled.empty_columns(n, transformed=True)
    batch = compiled.sample(n, 20240002, with_fingerprint=True, transformed_out=tb)
    assert bool(batch.mask[:n].all().item())
    assert bool(compiled.contains(batch).all().item())
    assert int(batch.n_failed.sum().item()) == 0
    back = compiled.reverse(tb)
    for a, b in zip(back.columns, batch.columns):
        assert torch.equal(a, b)
    tcols = tb.columns
    t = 0
    for s, (spec_kind, shape) in enumerate(zip(compiled.dim_kinds, compiled.dim_shapes)):
        if spec_kind == 'categorical':
            k = len(compiled.dim_categories[s])
            total = sum(tcols[t + j] for j in range(k))
            assert bool((total == 1.0).all().item())
            t += k
        else:
            t += 1
    again = compiled.sample(n, 20240002, with_fingerprint=True)
    assert torch.equal(again.fp[:n], batch.fp[:n])
    t2 = compiled.transform(again)
    for a, b in zip(t2.columns, tcols):
        assert torch.equal(a, b)
    # fingerprints of distinct candidates are distinct
    fp = batch.fp[:n]
    packed = fp[:, 0] ^ (fp[:, 1] * 0x9E3779B97F4A7C15 % (1 << 63))
    assert torch.unique(packed).numel() == n


def test_cat32x100_chunked_one_hot_round_trip():
    """BASELINE.json config 4: 32 x 100 categories, 10M points, chunks of 1M
    (a 1M chunk of one-hot is 25.6 GB): reverse(transform(x)) == x and every
    one-hot block has exactly one 1."""
    from orion_b200 import spaces
    from orion_b200.transformer import build_required_space
    space = spaces.cat32x100()
    compiled = build_required_space('real', space).compiled()
    chunk, chunks = 1_000_000, 10
    tb = compiled.empty_columns(chunk, transformed=True)
    back = compiled.empty_columns(chunk)
    assert compiled.trow_bytes == 25600 and compiled.row_bytes == 128
    for c in range(chunks):
        batch = compiled.sample(chunk, 20240004, base_index=c * chunk)
        compiled.transform(batch, out=tb)
        compiled.reverse(tb, out=back)
        assert torch.equal(back.buf, batch.buf) or all(torch.equal(a, b) for a, b in zip(back.columns, batch.columns))
        if c == 0:
            tcols = tb.columns
            ones = sum(tcols[j] for j in range(100))
            assert bool((ones == 1.0).all().item())
            idx = batch.columns[0].long()
            picked = torch.stack(tcols[:100], dim=1).gather(1, idx[:, None])[:, 0]
            assert bool((picked == 1.0).all().item())


def test_full_size_suggest_and_dedup_exact_keep_mask():
    """BASELINE.json config 5's per-GPU batch (1.25e7 candidates of mixed64)
    through the one-kernel suggest and the history dedup, with planted
    duplicates: the keep mask must equal the one derived from counter equality
    (candidates are a function of their Philox counter; continuous dimensions
    never collide otherwise), the staged kernels must produce the same bytes,
    and a second pass over the same counters must keep nothing."""
    from orion_b200 import spaces
    from orion_b200.history import TrialHistory
    from orion_b200.transformer import build_required_space
    compiled = build_required_space('real', spaces.mixed64()).compiled()
    n, n_hist, seed = 12_500_000, 25_000_000, 20240005
    hist_base = 1 << 40
    history = TrialHistory(compiled, capacity=n_hist + 2 * n)
    hb = compiled.empty_columns(n)
    for start in range(0, n_hist, n):
        m = min(n, n_hist - start)
        b = compiled.sample(m, seed, base_index=hist_base + start, with_mask=False, with_fingerprint=True,
                            out=hb if m == n else None)
        history.insert_fingerprints(b.fp, m)
    assert len(history) == n_hist
    g = torch.Generator(device='cuda')
    g.manual_seed(1)
    idx = torch.arange(n, dtype=torch.int64, device='cuda')
    r = torch.rand(n, generator=g, device='cuda')
    idx = torch.where(r < 0.10, hist_base + torch.randint(0, n_hist, (n,), generator=g, device='cuda'), idx)
    early = torch.randint(0, n // 100, (n,), generator=g, device='cuda')
    idx = torch.where((r >= 0.10) & (r < 0.11), early, idx)
    # expected verdicts from the counters alone: not a history counter, and the first row with this counter
    order = torch.argsort(idx, stable=True)
    s = idx[order]
    first_sorted = torch.ones(n, dtype=torch.bool, device='cuda')
    first_sorted[1:] = s[1:] != s[:-1]
    first = torch.empty(n, dtype=torch.bool, device='cuda')
    first[order] = first_sorted
    want = first & (idx < hist_base)

    batch, tb, back = compiled.suggest(n, seed, index=idx)
    assert bool(batch.mask[:n].all().item())
    assert torch.equal(back.buf, batch.buf)                      # 'real': the round trip is the identity, byte for byte
    keep = history.dedup_fingerprints(back.fp, n)
    assert torch.equal(keep[:n].bool(), want)
    assert len(history) == n_hist + int(want.sum().item())
    # the staged kernels write the same bytes
    tb2 = compiled.empty_columns(n, transformed=True)
    b2 = compiled.sample(n, seed, index=idx, transformed_out=tb2)
    assert torch.equal(b2.buf, batch.buf) and torch.equal(tb2.buf, tb.buf)
    del tb
    k2 = compiled.reverse(tb2, out=b2, with_fingerprint=True)    # (reverse may overwrite the sampled columns)
    assert torch.equal(k2.fp[:n], back.fp[:n]) and torch.equal(k2.buf, back.buf)
    # idempotence: everything is history now
    again = history.dedup_fingerprints(back.fp, n)
    assert int(again[:n].sum().item()) == 0
