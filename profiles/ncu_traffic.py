#!/usr/bin/env python3
"""profiles/r2_ncu_traffic.json from an ncu report: DRAM bytes (read + write) per launch of the hot kernels, tagged with
the sha of the CUDA sources they were built from, so that bench.py quotes `roofline.traffic` only for the same build.

    python profiles/ncu_traffic.py gpurun_out/prof.ncu-rep[,second.ncu-rep] [shape] [scale] [n_gpus]   (run here, where ncu -i works)"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rep = sys.argv[1]
    shape = sys.argv[2] if len(sys.argv) > 2 else 'venice'
    scale = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    n = int(sys.argv[4]) if len(sys.argv) > 4 else 1
    scale_of = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    acc = {}
    for one in rep.split(','):
        out = subprocess.run(['ncu', '-i', one, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hdr, units = rows[0], rows[1]
        ix = {h: i for i, h in enumerate(hdr)}
        collect(rows[2:], ix, units, scale_of, acc)
    finish(rep, shape, scale, n, acc)


def collect(rows, ix, units, scale_of, acc):
    for r in rows:
        name = r[ix['Kernel Name']]
        key = next((k for k in ('k_matvec32', 'k_matvec', 'k_linearize32', 'k_linearize', 'k_pcg_step') if k in name), None)
        if key is None:
            continue
        tot = 0.0
        for m in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
            tot += float(r[ix[m]].replace(',', '')) * scale_of[units[ix[m]]]
        dur = float(r[ix['gpu__time_duration.sum']].replace(',', ''))
        if (key in ('k_matvec32', 'k_matvec') and dur < 40.0) or (key == 'k_pcg_step' and dur < 10.0):
            continue                                    # a launch past convergence (exits at its first instruction)
        acc.setdefault(key, []).append(tot)


def finish(rep, shape, scale, n, acc):
    import bench
    rec = {'report': ','.join(os.path.basename(x) for x in rep.split(',')), 'source_sha': bench.source_sha(), 'shape': shape, 'scale': scale, 'n_gpus': n,
           'kernels': {k: sum(v) / len(v) for k, v in acc.items()}, 'launches': {k: len(v) for k, v in acc.items()}}
    with open(os.path.join(ROOT, 'profiles', 'r2_ncu_traffic.json'), 'w') as f:
        json.dump(rec, f, indent=1)
    print(json.dumps(rec))


if __name__ == '__main__':
    main()
