#!/usr/bin/env python3
"""Per-source-line stall samples from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`.
    ncu -i rep --page source --print-source cuda,sass --csv --kernel-name regex:K --launch-count 1 | python ncu_lines.py [N]"""
import csv
import sys

n = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rows = list(csv.reader(sys.stdin))
cur_file, hdr, out = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
        idx = {}
        for i, h in enumerate(hdr):
            idx.setdefault(h, i)
    elif hdr and r[0].isdigit() and len(r) > idx['# Samples']:
        def I(k):
            try:
                return int(r[idx[k]])
            except Exception:
                return 0
        out.append((I('# Samples'), cur_file, int(r[0]), r[1].strip()[:80], I('Instructions Executed'),
                    {k[6:]: I(k) for k in idx if k.startswith('stall_') and '(' not in k}))
tot = sum(o[0] for o in out) or 1
print('total samples', tot)
for smp, f, ln, src, ne, st in sorted(out, key=lambda o: -o[0])[:n]:
    top = ' '.join('%s=%d' % (k, v) for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3] if v * 8 > smp)
    print('%5.1f%% %6d x%-8d %s:%-4d %-80s | %s' % (100.0 * smp / tot, smp, ne, f, ln, src, top))
agg = {}
for o in out:
    for k, v in o[5].items():
        agg[k] = agg.get(k, 0) + v
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
