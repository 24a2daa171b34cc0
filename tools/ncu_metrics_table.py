"""Per-kernel table from an `ncu --metrics ... --csv` log of one forward (tools/run_forward_once.py):
launches, time, DRAM bytes per launch, achieved DRAM GB/s, tensor-pipe activity, issue rate.
usage: python tools/ncu_metrics_table.py gpurun_out/forward_metrics_v10.csv [--json out.json]"""
import csv
import json
import re
import sys
from collections import OrderedDict

path = sys.argv[1]
lines = [l for l in open(path, newline='') if l.startswith('"')]
per = OrderedDict()  # launch id -> dict
for r in csv.DictReader(lines):
    k = r['ID']
    d = per.setdefault(k, {'name': re.sub(r'\(.*$', '', r['Kernel Name']).replace('void ', '').replace('srb::', '')
                           .replace('<unnamed>::', '').replace('(anonymous namespace)::', '')})
    v = float(r['Metric Value'].replace(',', '') or 0)
    u = r['Metric Unit']
    scale = {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6, 'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1.0)
    d[r['Metric Name']] = v * scale
agg = OrderedDict()
for d in per.values():
    a = agg.setdefault(d['name'], {'n': 0, 'us': 0.0, 'rd': 0.0, 'wr': 0.0, 'l2': 0.0, 'tc': 0.0, 'ipc': 0.0, 'regs': 0})
    t = d.get('gpu__time_duration.sum', 0.0)
    a['n'] += 1
    a['us'] += t
    a['rd'] += d.get('dram__bytes_read.sum', 0.0)
    a['wr'] += d.get('dram__bytes_write.sum', 0.0)
    a['l2'] += d.get('lts__t_bytes.sum', 0.0)
    a['tc'] += d.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 0.0) * t
    a['ipc'] += d.get('smsp__issue_active.avg.per_cycle_active', 0.0) * t
    a['regs'] = int(d.get('launch__registers_per_thread', 0))
tot = sum(a['us'] for a in agg.values())
print('| kernel | launches | total ms | share | avg us | DRAM MB/launch (rd+wr) | DRAM GB/s | L2 MB/launch | tensor pipe active % | issue slots busy % | regs |')
print('|---|---|---|---|---|---|---|---|---|---|---|')
out = {}
for name, a in sorted(agg.items(), key=lambda kv: -kv[1]['us']):
    n, us = a['n'], a['us']
    mb = (a['rd'] + a['wr']) / n / 1e6
    print(f"| `{name}` | {n} | {us / 1e3:.3f} | {100 * us / tot:.1f}% | {us / n:.1f} | {a['rd'] / n / 1e6:.1f} + {a['wr'] / n / 1e6:.1f} "
          f"| {(a['rd'] + a['wr']) / us / 1e3:.0f} | {a['l2'] / n / 1e6:.0f} | {a['tc'] / us:.1f} | {100 * a['ipc'] / us:.0f} | {a['regs']} |")
    out[name] = {'launches': n, 'total_us': round(us, 1), 'dram_bytes_per_launch': round((a['rd'] + a['wr']) / n),
                 'tensor_pipe_active_pct': round(a['tc'] / us, 2)}
print(f'\ntotal {tot / 1e3:.2f} ms over {sum(a["n"] for a in agg.values())} launches (serialised, L2 flushed once before the forward)')
if '--json' in sys.argv:
    json.dump(out, open(sys.argv[sys.argv.index('--json') + 1], 'w'), indent=1)
