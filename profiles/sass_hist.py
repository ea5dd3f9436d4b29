#!/usr/bin/env python3
"""Opcode histogram of the hot kernels' SASS (cuobjdump -sass of libsfmba.so): which staging path (LDGSTS = cp.async,
UBLKCP = cp.async.bulk, LDG/STG), which math pipes (DFMA/DADD/DMUL, FFMA, DMMA) each kernel really uses.

    python profiles/sass_hist.py > profiles/r2_sass_hot_kernels.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'structure-from-motion_b200', 'libsfmba.so')
HOT = ['k_linearize', 'k_matvec32', 'k_matvecILb0ELb0', 'k_pcg_step', 'k_cam_finish', 'k_camsum64', 'k_backsub_points',
       'k_cost_tiles', 'k_chunk_build', 'k_dense_gemm', 'k_triangulate', 'k_pnp']


def main():
    out = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True).stdout
    funcs, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            cur = m.group(1)
            funcs[cur] = collections.Counter()
            continue
        m = re.match(r'\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)', line)
        if m and cur:
            funcs[cur][m.group(1).split('.')[0]] += 1
    demangle = lambda n: subprocess.run(['c++filt', n], capture_output=True, text=True).stdout.strip()  # noqa: E731
    print('SASS opcode histograms, sm_100a, %s' % os.path.relpath(LIB, ROOT))
    for name, cnt in funcs.items():
        if not any(h in name for h in HOT):
            continue
        total = sum(cnt.values())
        keys = ['UBLKCP', 'SYNCS', 'LDGSTS', 'LDG', 'STG', 'LDS', 'STS', 'DFMA', 'DADD', 'DMUL', 'FFMA', 'FADD', 'FMUL', 'MUFU',
                'DMMA', 'SHFL', 'BAR', 'ATOM', 'ATOMG', 'RED', 'LDL', 'STL', 'F2F', 'MEMBAR', 'ACQBULK', 'UTMALDG']
        print('\n%s\n  %d instructions; %s' % (demangle(name)[:150], total,
              '  '.join('%s %d' % (k, cnt[k]) for k in keys if cnt[k])))
    allc = collections.Counter()
    for c in funcs.values():
        allc.update(c)
    print('\nwhole library: UBLKCP %d  LDGSTS %d  DMMA %d  UTMALDG %d  UTCHMMA/UTCQMMA %d' %
          (allc['UBLKCP'], allc['LDGSTS'], allc['DMMA'], allc['UTMALDG'], allc['UTCHMMA'] + allc['UTCQMMA']))


if __name__ == '__main__':
    main()
