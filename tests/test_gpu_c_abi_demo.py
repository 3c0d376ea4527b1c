"""The C ABI used from plain C (examples/c_abi_demo.c: no Python, no torch in
the process): compiled here with nvcc, run, and compared bit for bit with the
Python host layer driving the same library."""
import os
import shutil
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which('nvcc') is None, reason='needs nvcc to build the C demo')
def test_c_program_gets_the_same_points_as_the_python_host(tmp_path):
    from orion_b200.space import Integer, Real, Space
    exe = str(tmp_path / 'c_abi_demo')
    subprocess.run(['nvcc', '-O2', os.path.join(ROOT, 'examples', 'c_abi_demo.c'), '-I', os.path.join(ROOT, 'include'),
                    '-L', os.path.join(ROOT, 'orion_b200'), '-lob2', '-o', exe], check=True, capture_output=True)
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, 'orion_b200') + ':' + os.environ.get('LD_LIBRARY_PATH', ''))
    n, seed = 37, 20240001
    out = subprocess.run([exe, str(n), str(seed)], check=True, capture_output=True, text=True, env=env, timeout=120).stdout
    lines = out.strip().splitlines()
    assert lines[-1] == 'host-entry identical'
    rows = [l.split() for l in lines[:-1]]
    assert len(rows) == n
    x = np.array([float.fromhex(r[1]) for r in rows])
    k = np.array([int(r[2]) for r in rows], dtype=np.int64)
    assert all(r[3] == '1' and r[4] == '1' for r in rows)           # fused mask and ob2_contains agree: in space
    space = Space()
    space.register(Real('x', 'uniform', -5, 15))
    space.register(Integer('k', 'uniform', -3, 9))
    cols, failed = space.compiled().sample_host(n, seed)
    assert not failed.any()
    assert np.array_equal(cols[0], x) and np.array_equal(cols[1], k)
