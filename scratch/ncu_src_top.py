"""Aggregate an `ncu --page source --csv` export: top SASS lines by stall samples with their dominant stall reasons.
usage: python scratch/ncu_src_top.py file.csv [N] [min_addr_index max_addr_index]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
h = rows[hi]; data = rows[hi + 1:]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
lo_i = int(sys.argv[3]) if len(sys.argv) > 3 else 0
hi_i = int(sys.argv[4]) if len(sys.argv) > 4 else len(data)
ci = {k: i for i, k in enumerate(h)}
stall_cols = [k for k in h if k.startswith('stall_') and 'Not Issued' not in k]
def f(x):
    try: return float(x)
    except: return 0.0
tot = sum(f(r[ci['# Samples']]) for r in data)
print('total samples', tot, 'lines', len(data))
sel = list(enumerate(data))[lo_i:hi_i]
agg = {}
for i, r in sel:
    for k in stall_cols:
        agg[k] = agg.get(k, 0) + f(r[ci[k]])
print('stall mix:', ', '.join(f'{k[6:]} {v:.0f}' for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
top = sorted(sel, key=lambda ir: -f(ir[1][ci['# Samples']]))[:N]
for i, r in sorted(top):
    st = sorted(((f(r[ci[k]]), k[6:]) for k in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {f(r[ci['# Samples']]):7.0f} {f(r[ci['Instructions Executed']]):9.0f}  {r[ci['Source']][:70]:70s} {st[0][1]}:{st[0][0]:.0f} {st[1][1]}:{st[1][0]:.0f}")
