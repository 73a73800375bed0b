"""In-tree build of libdegenprimer_b200.so (nvcc, sm_100a only).

`python -m degenprimer_b200.build` or `__graft_entry__.build()`.  nvcc
cross-compiles without a GPU; the .so is git-ignored but travels to the GPU
box with the repository snapshot.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libdegenprimer_b200.so')
SOURCES = ['dp_ctx.cu', 'dp_template.cu', 'dp_scan.cu', 'dp_dg.cu', 'dp_comm.cu', 'dp_jit.cu', 'dp_calib.cu', 'dp_static.cu', 'dp_mixture.cu']
NVCC_FLAGS = ['-std=c++17', '-O3', '-gencode', 'arch=compute_100a,code=sm_100a',
              '-lineinfo', '-Xcompiler', '-fPIC']


def _nvcc():
    exe = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(exe):
        raise RuntimeError('nvcc not found: cannot build libdegenprimer_b200.so')
    return exe


STAMP = LIB + '.sources'


def _source_digest():
    """Digest of everything the library is built from.  File times do not survive the copy to a GPU box
    (and eight ranks starting at once must not all decide to rebuild), so staleness is decided by content."""
    import hashlib
    h = hashlib.sha256()
    deps = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC))
    deps.append(os.path.join(os.path.dirname(HERE), 'include', 'degenprimer_b200.h'))
    for d in deps:
        h.update(os.path.basename(d).encode() + b'\0')
        with open(d, 'rb') as fh:
            h.update(fh.read())
    h.update(' '.join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _stale():
    if not os.path.exists(LIB) or not os.path.exists(STAMP):
        return True
    try:
        with open(STAMP) as fh:
            return fh.read().strip() != _source_digest()
    except OSError:
        return True


def _embed_device_header(objdir):
    """dp_device.cuh as a C++ raw string literal: the header of the run-time compiled kernels."""
    text = open(os.path.join(CSRC, 'dp_device.cuh')).read()
    assert ')DPJIT"' not in text
    out = os.path.join(objdir, 'dp_jit_embed.inc')
    body = 'R"DPJIT(' + text + ')DPJIT"\n'
    if not os.path.exists(out) or open(out).read() != body:
        with open(out, 'w') as f:
            f.write(body)


def build_native(force=False, verbose=False):
    """Compile every CUDA source into one shared library.  Returns its path."""
    if not force and not _stale():
        return LIB
    import fcntl
    objdir = os.path.join(HERE, 'build')
    os.makedirs(objdir, exist_ok=True)
    with open(os.path.join(objdir, '.lock'), 'w') as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)          # one builder at a time (ranks of one node share the tree)
        try:
            if not force and not _stale():        # somebody else built it while this process waited
                return LIB
            return _build_locked(objdir, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(objdir, verbose):
    nvcc = _nvcc()
    _embed_device_header(objdir)
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace('.cu', '.o'))
        cmd = [nvcc] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + \
              ['-c', os.path.join(CSRC, src), '-o', obj]
        procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    objs = []
    for src, obj, p in procs:
        out = p.communicate()[0].decode()
        if p.returncode:
            raise RuntimeError('nvcc failed on %s:\n%s' % (src, out))
        if verbose:
            sys.stderr.write(out)
        objs.append(obj)
    tmp = LIB + '.tmp'
    cmd = [nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', tmp] + objs + ['-ldl']
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode:
        raise RuntimeError('link failed:\n%s' % r.stdout.decode())
    os.replace(tmp, LIB)
    with open(STAMP + '.tmp', 'w') as fh:
        fh.write(_source_digest() + '\n')
    os.replace(STAMP + '.tmp', STAMP)
    return LIB


if __name__ == '__main__':
    print(build_native(force='--force' in sys.argv, verbose='-v' in sys.argv))
