"""Aggregates the warp-stall samples of an .ncu-rep per CUDA source line (needs -lineinfo builds and
`ncu --import-source on`).  usage: python tools/ncu_lines.py report.ncu-rep [top_n]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass,cuda'], capture_output=True,
                     text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
H = next(r for r in rows if '# Samples' in r)
si = H.index('# Samples')
names = [h for h in H if h.startswith('stall_') and 'Not Issued' not in h]
idx = {n: H.index(n) for n in names}
cur, src = '', []
for r in rows:
    if r and r[0] == 'File Path':
        cur = r[1].split('/')[-1]
    if len(r) > si and r[0].isdigit() and r[si].isdigit():
        src.append((cur, r))
tot = sum(int(r[si]) for _, r in src) or 1
print('total samples', tot)
for f, r in sorted(src, key=lambda fr: -int(fr[1][si]))[:top_n]:
    st = {n: int(r[idx[n]]) for n in names if r[idx[n]].isdigit() and int(r[idx[n]]) > 0}
    top = ', '.join(f'{k[6:]} {v}' for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(f'{f[:20]:20s}{r[0]:>5} {int(r[si]) / tot:6.3f}  [{top}]  {r[1].strip()[:80]}')
