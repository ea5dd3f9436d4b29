#!/usr/bin/env python3
"""Builds variants/libsfmba_<name>.so: the in-tree objects, with the named sources recompiled under extra -D flags.
A/B builds for profiles/ab_bench.py (SFMBA_LIB picks the library at run time).

    python profiles/build_variant.py <name> <source.cu>[,<source.cu>...] -DFLAG=VALUE ..."""
import importlib.util
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location('b', os.path.join(ROOT, 'structure-from-motion_b200', 'build.py'))
b = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b)


def main():
    name, srcs, defs = sys.argv[1], sys.argv[2].split(','), sys.argv[3:]
    b.build_library()
    out = os.path.join(ROOT, 'variants')
    os.makedirs(out, exist_ok=True)
    objs = []
    for src in b.SOURCES:
        o = os.path.join(b.OBJ, src.replace('.cu', '.o'))
        if src in srcs:
            o = os.path.join(out, '%s_%s' % (name, src.replace('.cu', '.o')))
            subprocess.run([b._nvcc()] + b.NVCC_FLAGS + defs + ['-c', os.path.join(b.CSRC, src), '-o', o], check=True)
        objs.append(o)
    lib = os.path.join(out, 'libsfmba_%s.so' % name)
    subprocess.run([b._nvcc(), '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', lib] + objs +
                   ['-Xcompiler', '-fPIC', '-lcudart_static', '-ldl', '-lpthread', '-lrt'], check=True)
    print(lib)


if __name__ == '__main__':
    main()
