"""Short driver for ncu captures of the sampler variants (run under gpurun):
mixed64 plain sample, sample + fingerprint (config 5's sampler), one-kernel suggest,
real16 sample at 1e8; `reps` launches each, CUDA-event times printed as JSON."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from orion_b200 import spaces  # noqa: E402
from orion_b200.transformer import build_required_space  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 10
n = 12_500_000
out = {"rounds": rounds}


def timed(name, fn):
    ts = []
    for r in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(r)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    out[name] = min(ts)


space = spaces.mixed64()
plain = space.compiled(philox_rounds=rounds)
buf = plain.empty_columns(n)
timed('mixed64 sample', lambda r: plain.sample(n, 7, base_index=r * n, with_mask=False, out=buf))
timed('mixed64 sample+mask', lambda r: plain.sample(n, 7, base_index=r * n, with_mask=True, out=buf))
timed('mixed64 sample+fingerprint', lambda r: plain.sample(n, 7, base_index=r * n, with_mask=True, with_fingerprint=True, out=buf))
tc = build_required_space('real', space).compiled(philox_rounds=rounds)
tb = tc.empty_columns(n, transformed=True)
bk = tc.empty_columns(n)
timed('mixed64 suggest', lambda r: tc.suggest(n, 7, base_index=r * n, out=buf, transformed_out=tb, back_out=bk))
timed('mixed64 suggest back==cols', lambda r: tc.suggest(n, 7, base_index=r * n, out=buf, transformed_out=tb, back_out=buf))
timed('mixed64 sample+transform', lambda r: tc.sample(n, 7, base_index=r * n, with_mask=True, out=buf, transformed_out=tb))
del tb, bk, buf
r16 = spaces.real16().compiled(philox_rounds=rounds)
m = 100_000_000
b16 = r16.empty_columns(m)
timed('real16 sample 1e8', lambda r: r16.sample(m, 7, base_index=r * m, with_mask=False, out=b16))
print(json.dumps(out))
