"""Build libstransfer_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m styletransfer_b200.build [--force] [--verbose]

Every ``csrc/*.cu`` is compiled to its own object (in parallel, cached under ``build/`` by a digest
of the source, the headers and the flags) and the objects are linked into one shared library.
``--use_fast_math`` is a PER-FILE choice: only the tensor-core convolution kernels (whose epilogues are
adds, compares and bf16/fp16 conversions) get it; the normalisation / loss / optimizer files are
compiled with IEEE division and square root.
"""
import concurrent.futures
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
OBJ_DIR = os.path.join(HERE, 'build')
# source -> extra flags
FAST = ['--use_fast_math']
SOURCES = {
    'api.cu': [], 'conv_tc.cu': FAST, 'conv_tcx.cu': FAST, 'conv_imgrad_tc.cu': FAST, 'conv_generic.cu': FAST, 'conv_image.cu': FAST,
    'conv9.cu': FAST, 'conv9_tc.cu': FAST, 'contract.cu': FAST, 'wgrad_tc.cu': FAST,
    'elementwise.cu': [], 'norm_losses.cu': [], 'optim.cu': [], 'no_alloc.cu': [],
}
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
              '-Xcompiler', '-fPIC', '-Xptxas', '-v']
if os.environ.get('STRANSFER_CONV_TRACE') == '1':       # per-tile timeline of conv_tc (tools/trace_conv.py), never for product builds
    NVCC_FLAGS = NVCC_FLAGS + ['-DST_CONV_TRACE']
LINK_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-shared', '-cudart', 'shared']


def _nvcc():
    for cand in (shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found')


def _headers_digest():
    h = hashlib.sha256()
    files = [os.path.join(CSRC, n) for n in sorted(os.listdir(CSRC)) if n.endswith(('.cuh', '.h'))]
    files.append(os.path.join(INCLUDE, 'stransfer_b200.h'))
    for path in files:
        with open(path, 'rb') as f:
            h.update(os.path.basename(path).encode())      # never the absolute path: the tree is copied to other roots
            h.update(f.read())
    return h.hexdigest()


def _source_digest(name, hdr):
    h = hashlib.sha256()
    with open(os.path.join(CSRC, name), 'rb') as f:
        h.update(name.encode())
        h.update(f.read())
    h.update(hdr.encode())
    h.update(' '.join(NVCC_FLAGS + SOURCES[name]).encode())
    return h.hexdigest()


CONV_TC_SOURCES = ('conv_tc.cu', 'conv_tc_epilogue.cuh', 'ptx.cuh', 'common.cuh')


def kernel_sources_digest():
    """Digest of the sources of the dominant kernel (conv_tc: kernel, epilogue, PTX wrappers, common helpers): ties a
    committed ncu capture to the kernel it profiled (bench.py reads profiles/conv_tc_traffic.json only when this matches)."""
    h = hashlib.sha256()
    for name in CONV_TC_SOURCES:
        with open(os.path.join(CSRC, name), 'rb') as f:
            h.update(name.encode())
            h.update(f.read())
    return h.hexdigest()[:16]


def _digest():
    hdr = _headers_digest()
    h = hashlib.sha256()
    for name in sorted(SOURCES):
        h.update(_source_digest(name, hdr).encode())
    h.update(' '.join(LINK_FLAGS).encode())
    return h.hexdigest()


def up_to_date():
    if not (os.path.exists(LIB) and os.path.exists(STAMP)):
        return False
    with open(STAMP) as f:
        return f.read().strip() == _digest()


def _compile_one(nvcc, name, digest, force):
    obj = os.path.join(OBJ_DIR, name + '.o')
    tag = obj + '.digest'
    if not force and os.path.exists(obj) and os.path.exists(tag):
        with open(tag) as f:
            if f.read().strip() == digest:
                return name, 0, '(cached)\n'
    cmd = [nvcc] + NVCC_FLAGS + SOURCES[name] + ['-I', INCLUDE, '-c', '-o', obj, os.path.join(CSRC, name)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode == 0:
        with open(tag, 'w') as f:
            f.write(digest)
    return name, res.returncode, ' '.join(cmd) + '\n' + res.stdout + res.stderr


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
            nvcc = _nvcc()
            os.makedirs(OBJ_DIR, exist_ok=True)
            hdr = _headers_digest()
            logs, failed = [], []
            with concurrent.futures.ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 4)) as pool:
                jobs = [pool.submit(_compile_one, nvcc, n, _source_digest(n, hdr), force) for n in sorted(SOURCES)]
                for j in jobs:
                    name, rc, log = j.result()
                    logs.append(f'==== {name}\n{log}')
                    if rc != 0:
                        failed.append(name)
            tmp = LIB + f'.tmp{os.getpid()}'
            if not failed:
                cmd = [nvcc] + LINK_FLAGS + ['-o', tmp] + [os.path.join(OBJ_DIR, n + '.o') for n in sorted(SOURCES)]
                res = subprocess.run(cmd, capture_output=True, text=True)
                logs.append('==== link\n' + ' '.join(cmd) + '\n' + res.stdout + res.stderr)
                if res.returncode != 0:
                    failed.append('link')
            log = '\n'.join(logs)
            with open(os.path.join(HERE, 'build.log'), 'w') as f:
                f.write(log)
            if failed:
                if os.path.exists(tmp):
                    os.unlink(tmp)
                raise RuntimeError(f'nvcc failed ({failed}):\n' + log[-8000:])
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
