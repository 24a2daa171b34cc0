"""Digest of one `ncu --set full` report: headline metrics, stall reasons, hottest SASS lines.
usage: python tools/ncu_digest.py report.ncu-rep [n_hot_lines] [kernel-name regex when the report holds several kernels]"""
import csv, io, subprocess, sys

rep = sys.argv[1]
nhot = int(sys.argv[2]) if len(sys.argv) > 2 else 30
ksel = []
if len(sys.argv) > 3:  # a kernel-name regex, or '#N' = the N-th launch of the report
    ksel = ['--launch-skip', sys.argv[3][1:], '--launch-count', '1'] if sys.argv[3].startswith('#') else ['-k', 'regex:' + sys.argv[3]]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'] + ksel, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[-1]
d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
want = ['gpu__time_duration.sum', 'sm__cycles_elapsed.avg', 'launch__registers_per_thread', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.per_cycle_active', 'smsp__warps_active.avg.per_cycle_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores']
for k in want:
    if k in d:
        print(f'{k:75s} {d[k][0]:>16s} {d[k][1]}')
st = []
for h, (v, u) in d.items():
    if 'pcsamp_warps_issue_stalled' in h and 'not_issued' not in h:
        try:
            st.append((float(v), h.replace('smsp__pcsamp_warps_issue_stalled_', '')))
        except ValueError:
            pass
tot = sum(v for v, _ in st) or 1
print('stalls:', ', '.join(f'{h} {100 * v / tot:.1f}%' for v, h in sorted(st, reverse=True)[:8]))
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'] + ksel, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
his = [i for i, r in enumerate(rows) if 'Source' in r and 'Address' in r]  # one header row per kernel of the report
sec = int(sys.argv[3][1:]) if len(sys.argv) > 3 and sys.argv[3].startswith('#') and len(his) > 1 else 0
hi = his[sec]
hdr = rows[hi]; data = rows[hi + 1:his[sec + 1] if sec + 1 < len(his) else len(rows)]
data = [r for r in data if len(r) == len(hdr)]
ia, isamp = hdr.index('Source'), hdr.index('Warp Stall Sampling (All Samples)')
tot = sum(int(r[isamp]) for r in data if r[isamp].isdigit()) or 1
idx = sorted(range(len(data)), key=lambda i: -int(data[i][isamp] or 0))[:nhot]
print(f'hot SASS (of {tot} samples):')
for i in sorted(idx):
    print(f'{i:5d} {100 * int(data[i][isamp]) / tot:5.1f}%  {data[i][ia].strip()[:100]}')
