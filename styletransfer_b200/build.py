"""Build libstransfer_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m styletransfer_b200.build [--force] [--verbose]
"""
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
INCLUDE = os.path.join(os.path.dirname(HERE), 'include')
LIB = os.path.join(HERE, 'libstransfer_b200.so')
STAMP = os.path.join(HERE, '.build_stamp')
SOURCES = ['api.cu', 'conv_tc.cu', 'conv_tcx.cu', 'conv_generic.cu', 'conv_image.cu', 'conv9.cu', 'conv9_tc.cu', 'contract.cu', 'wgrad_tc.cu', 'elementwise.cu', 'norm_losses.cu', 'optim.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
              '--use_fast_math', '-Xcompiler', '-fPIC', '-Xptxas', '-v', '-shared', '-cudart', 'shared']


def _nvcc():
    for cand in (shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found')


def _digest():
    h = hashlib.sha256()
    files = sorted(os.listdir(CSRC)) + [os.path.join(INCLUDE, 'stransfer_b200.h')]
    for name in files:
        path = name if os.path.isabs(name) else os.path.join(CSRC, name)
        with open(path, 'rb') as f:
            h.update(os.path.basename(name).encode())      # never the absolute path: the tree is copied to other roots
            h.update(f.read())
    h.update(' '.join(NVCC_FLAGS).encode())
    return h.hexdigest()


def up_to_date():
    if not (os.path.exists(LIB) and os.path.exists(STAMP)):
        return False
    with open(STAMP) as f:
        return f.read().strip() == _digest()


def build(force=False, verbose=False):
    """Builds under an exclusive file lock and publishes the library with an atomic rename: the ranks of a
    torchrun launch may all get here at once; one of them compiles, the others wait and find it up to date,
    and no process can ever dlopen a half-written file."""
    import fcntl
    with open(os.path.join(HERE, '.build_lock'), 'w') as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and up_to_date():
                return LIB
            tmp = LIB + f'.tmp{os.getpid()}'
            cmd = [_nvcc()] + NVCC_FLAGS + ['-I', INCLUDE, '-o', tmp] + [os.path.join(CSRC, s) for s in SOURCES]
            res = subprocess.run(cmd, capture_output=True, text=True)
            log = res.stdout + res.stderr
            with open(os.path.join(HERE, 'build.log'), 'w') as f:
                f.write(' '.join(cmd) + '\n' + log)
            if res.returncode != 0:
                if os.path.exists(tmp):
                    os.unlink(tmp)
                raise RuntimeError('nvcc failed:\n' + log[-6000:])
            if verbose:
                print(log)
            os.replace(tmp, LIB)
            with open(STAMP + '.tmp', 'w') as f:
                f.write(_digest())
            os.replace(STAMP + '.tmp', STAMP)
            return LIB
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


if __name__ == '__main__':
    path = build(force='--force' in sys.argv, verbose='--verbose' in sys.argv)
    print('built', path)
