"""Phase-level breakdown of a kernel from an .ncu-rep: SASS is cut at every BAR.SYNC (and at the
mbarrier wait), and warp-instructions, stall samples and shared-memory wavefronts are summed per
segment.  Run here (no GPU needed).
    python tools/ncu_phases.py gpurun_out/prof.ncu-rep <kernel-regex> [launch index]"""
import csv
import subprocess
import sys

rep, kre = sys.argv[1], sys.argv[2]
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', f'regex:{kre}'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(r for r in rows if r and r[0] == 'Address')
# one block per captured launch; an optional third argument picks one of them
blocks, cur_b = [], None
for r in rows:
    if r and r[0] == 'Address':
        cur_b = []
        blocks.append(cur_b)
    elif cur_b is not None and len(r) == len(hdr) and r[0].startswith('0x'):
        cur_b.append(r)
data = blocks[int(sys.argv[3])] if len(sys.argv) > 3 else [r for b in blocks for r in b]
ix = {n: hdr.index(n) for n in hdr}
def num(r, n):
    try:
        return float(r[ix[n]])
    except Exception:
        return 0.0
names = ['# Samples', 'Instructions Executed', 'L1 Wavefronts Shared', 'L1 Wavefronts Shared Excessive',
         'stall_barrier', 'stall_long_sb', 'stall_short_sb', 'stall_mio', 'stall_not_selected', 'stall_wait', 'stall_math',
         'stall_lg', 'stall_selected', 'stall_no_inst', 'stall_dispatch', 'stall_branch_resolving']
segs = []
cur = {n: 0.0 for n in names}
cur['first'] = None
cur['ops'] = {}
def flush(tag):
    global cur
    cur['end'] = tag
    segs.append(cur)
    cur = {n: 0.0 for n in names}
    cur['first'] = None
    cur['ops'] = {}
for r in data:
    src = r[ix['Source']].strip()
    if cur['first'] is None:
        cur['first'] = src
    for n in names:
        cur[n] += num(r, n)
    op = src.split()[0] if not src.startswith('@') else src.split()[1]
    op = op.split('.')[0]
    cur['ops'][op] = cur['ops'].get(op, 0) + num(r, 'Instructions Executed')
    if 'BAR.SYNC' in src or 'SYNCS.PHASECHK' in src or src.startswith('EXIT'):
        flush(src)
flush('end')
ts = sum(s['# Samples'] for s in segs)
ti = sum(s['Instructions Executed'] for s in segs)
print(f'total samples {ts:.0f}, warp-instructions {ti:.0f}')
print(' seg  samp%  inst%   instr      shWf     shExc   | bar long short  mio nsel wait math  sel | top ops')
for i, s in enumerate(segs):
    if s['Instructions Executed'] == 0 and s['# Samples'] == 0:
        continue
    ops = sorted(s['ops'].items(), key=lambda kv: -kv[1])[:6]
    opstr = ' '.join(f'{k}:{100*v/max(s["Instructions Executed"],1):.0f}' for k, v in ops)
    print(f"{i:4d} {100*s['# Samples']/ts:6.1f} {100*s['Instructions Executed']/ti:6.1f} {s['Instructions Executed']:9.0f} "
          f"{s['L1 Wavefronts Shared']:9.0f} {s['L1 Wavefronts Shared Excessive']:9.0f} | "
          f"{s['stall_barrier']:4.0f} {s['stall_long_sb']:4.0f} {s['stall_short_sb']:4.0f} {s['stall_mio']:4.0f} "
          f"{s['stall_not_selected']:4.0f} {s['stall_wait']:4.0f} {s['stall_math']:4.0f} {s['stall_selected']:4.0f} | {opstr}")
