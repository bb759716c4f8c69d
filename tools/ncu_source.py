"""Per-source-line hot spots from an .ncu-rep (needs -lineinfo and --import-source on).
    python tools/ncu_source.py gpurun_out/prof.ncu-rep <kernel-regex> [topN]"""
import csv
import subprocess
import sys

rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda', '--kernel-name', f'regex:{kre}'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None
data = []
for r in rows:
    if r and r[0] in ('Line', '#', 'Address') or (hdr is None and 'Source' in r):
        hdr = r
        continue
    if hdr and len(r) >= len(hdr) - 2:
        data.append(r)
if not hdr:
    print(out[:2000])
    sys.exit(1)
def col(name):
    return hdr.index(name) if name in hdr else None
ci = {n: col(n) for n in ['Source', '# Samples', 'Instructions Executed', 'L1 Wavefronts Shared Excessive', 'L1 Wavefronts Shared',
                          'stall_barrier', 'stall_long_sb', 'stall_short_sb', 'stall_mio', 'stall_not_selected', 'stall_wait', 'stall_math', 'stall_lg']}
def num(r, n):
    i = ci[n]
    try:
        return float(r[i]) if i is not None and r[i] not in ('', '-') else 0.0
    except ValueError:
        return 0.0
tot_s = sum(num(r, '# Samples') for r in data)
tot_i = sum(num(r, 'Instructions Executed') for r in data)
print(f'total samples {tot_s:.0f}  total warp-instructions {tot_i:.0f}')
data.sort(key=lambda r: -num(r, '# Samples'))
print('samp%  inst%   shExc/shWf   bar  long short mio  notsel wait | source')
for r in data[:top]:
    print(f"{100*num(r,'# Samples')/max(tot_s,1):5.1f} {100*num(r,'Instructions Executed')/max(tot_i,1):5.1f}  "
          f"{num(r,'L1 Wavefronts Shared Excessive'):9.0f}/{num(r,'L1 Wavefronts Shared'):9.0f} "
          f"{num(r,'stall_barrier'):5.0f} {num(r,'stall_long_sb'):5.0f} {num(r,'stall_short_sb'):5.0f} {num(r,'stall_mio'):5.0f} "
          f"{num(r,'stall_not_selected'):5.0f} {num(r,'stall_wait'):5.0f} | {r[ci['Source']].strip()[:110]}")
