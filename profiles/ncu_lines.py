"""Stall samples and executed instructions per CUDA source line of an ncu report (needs -lineinfo + --import-source on).
usage: python profiles/ncu_lines.py report.ncu-rep [top_n]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'], capture_output=True,
                     text=True).stdout
cur, hdr, out = None, None, []
for r in csv.reader(txt.splitlines()):
    if not r:
        continue
    if r[0] == 'File Path':
        cur = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
    elif hdr and r[0].isdigit() and len(r) > 8 and r[6].isdigit():
        out.append((cur, r))
ismp, iex = hdr.index('# Samples'), hdr.index('Instructions Executed')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[ismp]) for _, r in out) or 1
tote = sum(int(r[iex]) for _, r in out) or 1
print(f'samples {tot}  warp-instructions {tote}')
for c, r in sorted(out, key=lambda cr: -int(cr[1][ismp]))[:top_n]:
    st = sorted(((int(r[i]), hdr[i][6:]) for i in stall_cols if r[i].isdigit() and int(r[i]) > 0), reverse=True)[:3]
    print(f'{c:14s} L{r[0]:>4s} smp {int(r[ismp]) / tot:6.3f} ins {int(r[iex]) / tote:6.3f}  '
          f'{" ".join(f"{n}:{v}" for v, n in st):40s} {r[1].strip()[:90]}')
