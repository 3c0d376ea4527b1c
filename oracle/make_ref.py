#!/usr/bin/env python
"""Stage the UNMODIFIED reference package as ``oracle/_ref/`` so that it travels
to the GPU box (test infrastructure -- nothing under ``orion_b200/`` imports it).

    python oracle/make_ref.py          # or: __graft_entry__.build()

The reference (DatCorno/orion) is ~4 k lines of pure Python with no build step:
"installing" it is copying ``src/orion`` -- the same files
``pip install --target oracle/_ref /root/reference`` would put there (pip itself
cannot run: ``setup.py`` imports versioneer and ``setup_requires`` pytest-runner,
and there is no network).  ``oracle/_ref/`` is git-ignored (no reference source
enters the history) but NOT gpurun-ignored, like the built ``libob2.so``.

Only runs where ``/root/reference`` exists (the build container); on the GPU box
the staged copy is used as it arrived.  The harness that imports it --
``oracle/ref_harness.py`` -- supplies the three accommodations SURVEY.md 8(c)
lists, none of which edits a reference file.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = '/root/reference/src/orion'
DST = os.path.join(HERE, '_ref')


def tree_digest(root):
    h = hashlib.sha256()
    n = 0
    for base, dirs, files in os.walk(root):
        dirs.sort()
        for name in sorted(files):
            if name.endswith('.py'):
                path = os.path.join(base, name)
                h.update(os.path.relpath(path, root).encode())
                h.update(open(path, 'rb').read())
                n += 1
    return h.hexdigest(), n


def stage(force=False):
    """Copy the package; returns the path of the staged tree, or None when
    there is no reference checkout here and nothing was staged before."""
    if not os.path.isdir(SRC):
        return DST if os.path.isdir(os.path.join(DST, 'orion')) else None
    digest, n_files = tree_digest(SRC)
    stamp = os.path.join(DST, 'STAGED.json')
    if not force and os.path.exists(stamp):
        try:
            if json.load(open(stamp)).get('sha256') == digest:
                return DST
        except ValueError:
            pass
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    os.makedirs(DST)
    shutil.copytree(SRC, os.path.join(DST, 'orion'), ignore=shutil.ignore_patterns('__pycache__', '*.pyc'))
    with open(stamp, 'w') as f:
        json.dump({"source": SRC, "sha256": digest, "python_files": n_files,
                   "how": "shutil.copytree of the reference's src/orion, byte for byte (oracle/make_ref.py)"}, f)
    return DST


if __name__ == '__main__':
    out = stage(force='--force' in sys.argv)
    print(out if out else "no reference checkout at %s and nothing staged" % SRC)
