#!/usr/bin/env python3
"""Builds degenprimer_b200/libvariant_<name>.so: the library with extra -D flags (kernel geometry experiments;
DEGENPRIMER_B200_LIB=<path> makes _native load it).   python tools/build_variant.py V2 -DDP_DUPLEX_WARPS=2"""
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from degenprimer_b200 import build as b          # noqa: E402


def main():
    name, flags = sys.argv[1], sys.argv[2:]
    objdir = os.path.join(b.HERE, 'build', 'variant_' + name)
    os.makedirs(objdir, exist_ok=True)
    b.build_native()                                  # the embedded header of the run-time kernels
    nvcc = b._nvcc()
    procs = []
    for src in b.SOURCES:
        obj = os.path.join(objdir, src.replace('.cu', '.o'))
        procs.append((src, obj, subprocess.Popen([nvcc] + b.NVCC_FLAGS + flags + ['-c', os.path.join(b.CSRC, src), '-o', obj],
                                                  stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    objs = []
    for src, obj, p in procs:
        out = p.communicate()[0].decode()
        if p.returncode:
            raise SystemExit('nvcc failed on %s:\n%s' % (src, out))
        objs.append(obj)
    lib = os.path.join(b.HERE, 'libvariant_%s.so' % name)
    subprocess.check_call([nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', lib] + objs + ['-ldl'])
    print(lib)


if __name__ == '__main__':
    main()
