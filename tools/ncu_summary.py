"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel markdown table.
usage: python tools/ncu_summary.py launches.csv [skip_first_n]"""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rows = []
    with open(path, newline='') as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        v = float(r['Metric Value'].replace(',', ''))
        u = r['Metric Unit']
        ns = v * {'ns': 1, 'us': 1e3, 'ms': 1e6, 's': 1e9}.get(u, 1)
        name = r['Kernel Name']
        name = re.sub(r'\(.*$', '', name)
        rows.append((name, ns, r['Grid Size'], r['Block Size']))
    rows = rows[skip:]
    agg = OrderedDict()
    for name, ns, g, b in rows:
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ns
    tot = sum(a[1] for a in agg.values())
    print('| kernel | launches | total ms | share | avg us |')
    print('|---|---|---|---|---|')
    for name, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f'| `{name}` | {n} | {ns / 1e6:.3f} | {100 * ns / tot:.1f}% | {ns / n / 1e3:.1f} |')
    print(f'\ntotal {tot / 1e6:.2f} ms over {len(rows)} launches')


if __name__ == '__main__':
    main()
