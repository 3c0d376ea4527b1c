#!/usr/bin/env python
"""What limits the fused sampler?  Time it with pieces switched off (no
fingerprint / no transform / no mask) on mixed64, and a write-only memset of the
same footprint for the store ceiling."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from orion_b200 import spaces  # noqa: E402
from orion_b200.transformer import build_required_space  # noqa: E402


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    n = 12_500_000
    c = build_required_space('real', spaces.mixed64()).compiled()
    b = c.empty_columns(n)
    tb = c.empty_columns(n, transformed=True)
    out = {}
    out['sample+fp+transform'] = timed(lambda: c.sample(n, 1, with_mask=True, with_fingerprint=True, out=b, transformed_out=tb))
    out['sample+transform (no fp)'] = timed(lambda: c.sample(n, 1, with_mask=True, with_fingerprint=False, out=b, transformed_out=tb))
    out['sample+fp (no transform)'] = timed(lambda: c.sample(n, 1, with_mask=True, with_fingerprint=True, out=b))
    out['sample only'] = timed(lambda: c.sample(n, 1, with_mask=True, with_fingerprint=False, out=b))
    out['transform only'] = timed(lambda: c.transform(b, out=tb))
    both = torch.empty(b.buf.numel() + tb.buf.numel(), dtype=torch.uint8, device='cuda')
    out['memset of cols+tcols (%.1f GB)' % (both.numel() / 1e9)] = timed(lambda: both.zero_())
    out['copy tcols->tcols (%.1f GB r + w)' % (tb.buf.numel() / 1e9)] = timed(lambda: both[:tb.buf.numel()].copy_(tb.buf))
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
