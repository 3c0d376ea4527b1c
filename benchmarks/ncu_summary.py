#!/usr/bin/env python
"""Read an .ncu-rep (here, on the CPU box) and print the few metrics the profile notes quote per
kernel; with --traffic also rewrite profiles/traffic.json, stamped with the hash of the kernel
sources the capture was taken from (bench.py accepts the file only while that hash matches).

    python benchmarks/ncu_summary.py gpurun_out/prof_r2p.ncu-rep [--traffic "capture note"]
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SCALE = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1, 'ms': 1e-3, 'us': 1e-6, 'ns': 1e-9, 's': 1}


def main():
    rep = sys.argv[1]
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}

    def val(r, k):
        return float(r[idx[k]]) * SCALE.get(units[idx[k]], 1)

    out = []
    for r in rows[2:]:
        stalls = sorted(((float(r[i]), h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''))
                         for h, i in idx.items() if 'smsp__average_warps_issue_stalled' in h
                         and h.endswith('per_issue_active.ratio') and r[i] not in ('', 'n/a')), reverse=True)[:4]
        out.append({'kernel': r[idx['Kernel Name']].replace('void <unnamed>::', '').split('(<unnamed>')[0].split('(')[0],
                    'time_ms': val(r, 'gpu__time_duration.sum') * 1e3,
                    'dram_read_GB': val(r, 'dram__bytes_read.sum') / 1e9, 'dram_write_GB': val(r, 'dram__bytes_write.sum') / 1e9,
                    'issue_active_pct': float(r[idx['smsp__issue_active.avg.pct_of_peak_sustained_active']]),
                    'warps_active_pct': float(r[idx['sm__warps_active.avg.pct_of_peak_sustained_active']]),
                    'fp64_pipe_pct': float(r[idx['sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active']]),
                    'regs': int(float(r[idx['launch__registers_per_thread']])),
                    'warp_inst': float(r[idx['smsp__inst_executed.sum']]),
                    'constant_cache_hit_pct': float(r[idx['idc__request_hit_rate.pct']]) if 'idc__request_hit_rate.pct' in idx else None,
                    'top_stalls': ['%s %.2f' % (h, v) for v, h in stalls]})
    for o in out:
        print(json.dumps(o))
    if '--traffic' in sys.argv:
        import bench
        note = sys.argv[sys.argv.index('--traffic') + 1]
        tr = {'csrc_sha16': bench.csrc_sha16(), 'capture': note, 'detail': out}
        for o in out:
            key = 'suggest' if 'sample_kernel' in o['kernel'] else ('dedup' if 'set_dedup' in o['kernel'] else None)
            if key:
                tr[key] = tr.get(key, 0) + (o['dram_read_GB'] + o['dram_write_GB']) * 1e9
        json.dump(tr, open(os.path.join(ROOT, 'profiles', 'traffic.json'), 'w'), indent=1)
        print('profiles/traffic.json <-', tr['csrc_sha16'], {k: tr[k] / 1e9 for k in ('suggest', 'dedup') if k in tr})


if __name__ == '__main__':
    main()
