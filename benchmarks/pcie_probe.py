"""What the box's PCIe link gives a plain pinned device->host copy (the e2e leg's ceiling)."""
import torch
n = 3_200_000_000
d = torch.empty(n, dtype=torch.uint8, device='cuda')
h = torch.empty(n, dtype=torch.uint8).pin_memory()
for chunks in (1, 64):
    step = n // chunks
    for rep in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for c in range(chunks):
            h[c * step:(c + 1) * step].copy_(d[c * step:(c + 1) * step], non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
    print('%d copies of %.0f MB: %.1f GB/s' % (chunks, step / 1e6, n / e0.elapsed_time(e1) / 1e6))
