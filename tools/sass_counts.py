"""Per-kernel SASS mnemonic counts of libsrb.so (static instruction counts, `cuobjdump -sass`), written as a markdown table.
usage: python tools/sass_counts.py [libsrb.so] > profiles/rNN_sass_counts.md
Columns: UTC*MMA (tcgen05.mma), UTCBAR (tcgen05.commit), LDTM / STTM (tcgen05.ld / .st), UTMALDG / UTMASTG (TMA tensor
load / store), LDGSTS (cp.async), SYNCS (mbarrier), HMMA (mma.sync: the non-tcgen05 tensor path), FFMA (CUDA-core MACs)."""
import os
import re
import subprocess
import sys
from collections import OrderedDict

COLS = OrderedDict([('UTC*MMA', r'\bUTC[A-Z]*MMA'), ('UTCBAR', r'\bUTCBAR'), ('LDTM', r'\bLDTM'), ('STTM', r'\bSTTM'),
                    ('UTMALDG', r'\bUTMALDG'), ('UTMASTG', r'\bUTMASTG'), ('LDGSTS', r'\bLDGSTS'), ('SYNCS', r'\bSYNCS'),
                    ('HMMA', r'\bHMMA'), ('FFMA', r'\bFFMA')])


def main():
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(here, 'superresolution_b200', 'libsrb.so')
    sass = subprocess.run(['cuobjdump', '-sass', so], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(['c++filt'], input='\n'.join(re.findall(r'Function : (\S+)', sass)), capture_output=True,
                           text=True).stdout.splitlines()
    parts = re.split(r'Function : \S+', sass)[1:]
    rows = []
    for name, body in zip(names, parts):
        name = re.sub(r'\(.*$', '', name.replace('(anonymous namespace)::', '')).replace('srb::', '').replace('void ', '')
        n_inst = len(re.findall(r'^\s+/\*[0-9a-f]{4}\*/', body, flags=re.M))
        rows.append((name, n_inst, [len(re.findall(p, body)) for p in COLS.values()]))
    rows.sort(key=lambda r: (-(r[2][0] > 0), -(r[2][8] > 0), r[0]))
    print('| kernel | SASS instrs | ' + ' | '.join(COLS) + ' |')
    print('|---|---|' + '---|' * len(COLS))
    for name, n, c in rows:
        print(f'| `{name}` | {n} | ' + ' | '.join(str(v) if v else '·' for v in c) + ' |')


if __name__ == '__main__':
    main()
