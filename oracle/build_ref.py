"""Stage the UNMODIFIED Python reference under ``oracle/_ref/`` so that it travels to the GPU box.
TEST INFRASTRUCTURE ONLY.

The reference is pure Python: "building" it is byte-compiling ``/root/reference/stransfer/*.py`` where they lie
(``py_compile``, optimisation level 0) into sourceless modules ``oracle/_ref/stransfer/*.refbc`` (ordinary .pyc content; the
gpurun snapshot drops ``*.pyc`` files, hence the extension, and ``oracle/reference_shim.py`` installs a finder for it) -- no reference source
is copied anywhere, a SHA-256 manifest of the compiled sources is written beside the bytecode.  ``oracle/_ref/`` is
git-ignored but not gpurun-ignored (like the in-tree ``.so``), and the GPU box runs the same image (same CPython,
same bytecode magic), so ``bench.py --impl reference`` and
``cpu_baseline`` can time the reference's OWN code on the GPU box's host cores (``kind: "reference"``), imported
through ``oracle/reference_shim.py`` (stub modules for tensorboardX / imageio / matplotlib / colored_traceback,
``padding_mode='reflection'`` -> zeros as under the pinned torch 1.1.0, seeded random-init VGG-19).

    python -m oracle.build_ref          # in the build container (needs /root/reference)
"""
import hashlib
import json
import os
import py_compile
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get('STRANSFER_REFERENCE_SRC', '/root/reference')
DST = os.path.join(ROOT, 'oracle', '_ref')
EXT = '.refbc'


def staged():
    return os.path.isfile(os.path.join(DST, 'stransfer', 'network' + EXT))


def build():
    """Copies the reference package; returns the staging directory (or None when /root/reference is absent and
    nothing is staged)."""
    src_pkg = os.path.join(SRC, 'stransfer')
    if not os.path.isdir(src_pkg):
        return DST if staged() else None
    dst_pkg = os.path.join(DST, 'stransfer')
    shutil.rmtree(DST, ignore_errors=True)
    manifest = {}
    for base, _dirs, files in os.walk(src_pkg):
        for name in files:
            if not name.endswith('.py'):
                continue
            path = os.path.join(base, name)
            rel = os.path.relpath(path, SRC)
            out = os.path.join(DST, rel[:-3] + EXT)
            os.makedirs(os.path.dirname(out), exist_ok=True)
            py_compile.compile(path, cfile=out, dfile=rel, doraise=True, optimize=0)
            with open(path, 'rb') as f:
                manifest[rel] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(DST, 'MANIFEST.json'), 'w') as f:
        json.dump({'source': SRC, 'python': sys.version.split()[0], 'files': manifest}, f, indent=1, sort_keys=True)
    assert os.path.isfile(os.path.join(dst_pkg, 'network' + EXT))
    return DST


if __name__ == '__main__':
    out = build()
    print('reference staged at', out) if out else sys.exit('no /root/reference and nothing staged')
