#!/usr/bin/env python3
"""A/B of several builds of libsfmba.so on one box: runs bench.py once per library (SFMBA_LIB) and prints one row each.

    python profiles/ab_bench.py [--steps 10] [--extra "--shape venice"] name=path/to/lib.so ...

`name=default` stands for the in-tree build; `name=lib.so:VAR=VALUE[:VAR=VALUE]` also sets environment switches.  Rows go to stdout as JSON lines (kept under profiles/ when quoted)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    args = sys.argv[1:]
    steps, extra, reps = '10', [], 1
    libs = []
    while args:
        a = args.pop(0)
        if a == '--steps':
            steps = args.pop(0)
        elif a == '--extra':
            extra = args.pop(0).split()
        elif a == '--reps':
            reps = int(args.pop(0))
        else:
            libs.append(a.split('=', 1))
    for _ in range(reps):
        for name, path in libs:
            env = dict(os.environ)
            path, *sets = path.split(':')
            for kv in sets:
                env[kv.split('=', 1)[0]] = kv.split('=', 1)[1]
            if path != 'default':
                env['SFMBA_LIB'] = os.path.join(ROOT, path)
            r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--no-cpu', '--no-extras', '--steps', steps, '--warmup', '3'] + extra,
                               capture_output=True, text=True, env=env)
            if r.returncode != 0:
                print(json.dumps({'build': name, 'error': r.stderr[-600:]}), flush=True)
                continue
            j = json.loads(r.stdout.strip().splitlines()[-1])
            row = {'build': name, 'value': round(j['value'], 1), 'ms_per_step': round(j['ms_per_step'], 4),
                   'e2e': round(j['e2e']['value'], 1), 'e2e_ms': round(j['e2e']['ms_per_step'], 3),
                   'kernel_ms': {k: round(v, 4) for k, v in j['kernel_ms'].items()},
                   'phase_ms': {k: round(v, 4) for k, v in j['phase_ms_per_step'].items()},
                   'pcg_per_lm': j['config']['pcg_iterations_per_lm_iteration'], 'final_cost': j['final_cost'],
                   'launches': j['gpu_launches'], 'sm_mhz': j['clocks']['sm_mhz']}
            print(json.dumps(row), flush=True)


if __name__ == '__main__':
    main()
