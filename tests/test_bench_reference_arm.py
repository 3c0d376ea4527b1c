"""`bench.py --impl reference` (the CPU arm the driver runs beside the CUDA
arm): prints ONE JSON line with the contract's keys; under torchrun only rank 0
works.  Tiny samples of both CPU implementations -- the unmodified reference
package staged in oracle/_ref (when present) and the numpy port -- no GPU involved."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None, *flags):
    env = dict(os.environ)
    env.update(extra_env or {})
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '1',
                          '--warmup', '0', '--ref-cands', '1500', '--ref-literal-cands', '8'] + list(flags),
                         env=env, capture_output=True, text=True, timeout=300, check=True)
    return out.stdout


def _check_line(flags, kind):
    lines = [l for l in _run(None, *flags).splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['unit'] == 'candidates/s' and d['higher_is_better'] is True
    assert d['metric'].startswith('deduplicated candidates/sec') and d['value'] > 0 and d['vs_baseline'] is None
    assert d['steps'] == 1 and d['warmup'] == 0 and d['n_gpus'] == 1 and d['ms_per_step'] > 0
    assert d['cpu_baseline']['kind'] == kind and d['cpu_baseline']['cores'] >= 1
    assert d['cpu_baseline']['value'] == d['value'] and 'sample' in d['cpu_baseline']
    assert d['e2e'] == {'value': d['value'], 'unit': d['unit'], 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}
    assert 'workload' in d['config'] and 'model' not in d['config']


def test_reference_arm_prints_the_contract_line_for_the_numpy_port():
    _check_line(['--ref-impl', 'port'], 'port')


def test_reference_arm_runs_the_unmodified_reference_when_it_is_staged():
    from oracle import ref_harness
    if not ref_harness.available():
        import pytest
        pytest.skip("no reference package staged (oracle/make_ref.py)")
    _check_line([], 'reference')                       # auto picks oracle/_ref


def test_reference_arm_other_ranks_do_no_work():
    assert _run({'RANK': '1', 'WORLD_SIZE': '2', 'LOCAL_RANK': '1'}, '--gpus', '2').strip() == ''
