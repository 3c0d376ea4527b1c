"""Launches tests/multigpu_verify.py (exact keep masks of the sharded dedup on
real GPUs, peer-memory and NCCL exchanges) when the box has at least two GPUs;
on a single-GPU box it runs the same script with one rank, which still checks
the counter-derived mask over three consecutive steps."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(900)
def test_sharded_dedup_keep_masks_exact_on_the_visible_gpus(tmp_path):
    n_gpus = min(torch.cuda.device_count(), 2)
    out = str(tmp_path / 'verify.json')
    script = os.path.join(ROOT, 'tests', 'multigpu_verify.py')
    common = ['--cands', '2000000', '--history', '4000000', '--steps', '3', '--out', out]
    if n_gpus >= 2:
        cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(n_gpus),
               '--master-addr', '127.0.0.1', '--master-port', '29631', script] + common
    else:
        cmd = [sys.executable, script] + common
    proc = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=850)
    assert proc.returncode == 0, proc.stdout.decode('utf-8', 'replace')[-4000:]
    report = json.load(open(out))
    assert report['keep_mask_exact'] is True
    assert all(r['exact'] and r['mismatches'] == 0 for r in report['runs'])
    assert len(report['runs']) >= 3
