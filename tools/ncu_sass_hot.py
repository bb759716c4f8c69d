"""Top SASS instructions by stall samples for kernels matching a regex in an .ncu-rep (needs --set full).
    python tools/ncu_sass_hot.py rep.ncu-rep <kernel-substring> [topN] [which-match]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
blocks, cur = [], None
for r in rows:
    if r and r[0] == 'Kernel Name':
        cur = {'name': r[1], 'hdr': None, 'rows': []}
        blocks.append(cur)
    elif cur is not None and r and r[0] == 'Address':
        cur['hdr'] = r
    elif cur is not None and cur['hdr'] and len(r) == len(cur['hdr']):
        cur['rows'].append(r)
sel = [b for b in blocks if pat in b['name']]
b = sel[which]
h = b['hdr']
ix = {n: i for i, n in enumerate(h)}
stall_cols = [n for n in h if n.startswith('stall_')]
def num(r, n):
    try: return float(r[ix[n]])
    except Exception: return 0.0
tot = sum(num(r, '# Samples') for r in b['rows'])
print(b['name'][:150]); print('total samples', tot, 'instructions', len(b['rows']))
agg = {n: sum(num(r, n) for r in b['rows']) for n in stall_cols}
print('stall totals:', {k: int(v) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]})
order = sorted(range(len(b['rows'])), key=lambda i: -num(b['rows'][i], '# Samples'))[:top]
for i in sorted(order):
    r = b['rows'][i]
    st = sorted(((num(r, n), n[6:]) for n in stall_cols), reverse=True)[:3]
    print(f"{i:5d} {100*num(r,'# Samples')/tot:5.2f}%  exec={int(num(r,'Instructions Executed')):8d}  {r[ix['Source']].strip()[:70]:70s} " + ' '.join(f'{n}:{int(v)}' for v, n in st if v > 0))
