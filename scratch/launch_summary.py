"""Summarise an ncu launch list (csv with gpu__time_duration.sum + dram bytes per launch) of `bench.py --steps 2 --warmup 1`.
Takes the launches of ONE generate() (the last complete program pair before the per-launch timing pass) and prints per-kernel
shares; writes profiles/<tag>_traffic.json with the DRAM bytes per kernel function per step.
usage: python scratch/launch_summary.py gpurun_out/r1j_launches.csv r1j"""
import csv, json, sys, collections
rows = list(csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"')))
tag = sys.argv[2]
launch = collections.OrderedDict()
for r in rows:
    e = launch.setdefault(int(r['ID']), {'name': r['Kernel Name'].split('(')[0].split('::')[-1], 'grid': r['Grid Size'], 'block': r['Block Size']})
    v = float(r['Metric Value'].replace(',', ''))
    unit = r['Metric Unit']
    if r['Metric Name'].startswith('gpu__time'):
        e['us'] = v / 1e3 if unit in ('ns', 'nsecond') else (v if unit.startswith('us') else v * 1e3)
    else:
        scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[unit]
        e[r['Metric Name'].split('.')[0]] = v * scale
ls = list(launch.values())
# one generate() = the launches from one embed_kernel to the next
is_embed = lambda e: e['name'].startswith('embed')          # embed_kernel / embed_batch_kernel open a generate()
starts = [i for i, e in enumerate(ls) if is_embed(e) and (i == 0 or not is_embed(ls[i - 1]))]
seg = ls[starts[1]:starts[2]] if len(starts) > 2 else ls[starts[0]:]
tot = sum(e['us'] for e in seg)
print(f'# one generate(): {len(seg)} launches, sum of cold-cache serialised kernel times {tot:.1f} us')
agg = collections.OrderedDict()
for e in seg:
    a = agg.setdefault(e['name'], {'n': 0, 'us': 0.0, 'rd': 0.0, 'wr': 0.0})
    a['n'] += 1; a['us'] += e['us']; a['rd'] += e.get('dram__bytes_read', 0); a['wr'] += e.get('dram__bytes_write', 0)
print(f"{'kernel':28s} {'launches':>8s} {'us':>9s} {'share':>7s} {'dram rd MB':>11s} {'dram wr MB':>11s}")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]['us']):
    print(f"{k:28s} {a['n']:8d} {a['us']:9.1f} {a['us'] / tot:7.1%} {a['rd'] / 1e6:11.1f} {a['wr'] / 1e6:11.1f}")
print('# per launch:')
for i, e in enumerate(seg):
    print(f"{i:3d} {e['name']:26s} grid {e['grid']:>14s} {e['us']:8.1f} us  rd {e.get('dram__bytes_read', 0) / 1e6:8.1f} MB  wr {e.get('dram__bytes_write', 0) / 1e6:8.1f} MB")
out = {k: {'launches_per_step': a['n'], 'dram_bytes_per_step': a['rd'] + a['wr'], 'us_serialised': a['us']}
       for k, a in agg.items()}
# the CTA-pair form of the conv GEMM is the same kernel body: bench.py reports the family under conv_gemm_tc_kernel
if 'conv_gemm_tc_pair_kernel' in out:
    pr = out.pop('conv_gemm_tc_pair_kernel')
    base = out.setdefault('conv_gemm_tc_kernel', {'launches_per_step': 0, 'dram_bytes_per_step': 0.0, 'us_serialised': 0.0})
    for k in ('launches_per_step', 'dram_bytes_per_step', 'us_serialised'):
        base[k] += pr[k]
    base['pair_launches_per_step'] = pr['launches_per_step']
# the fused CBHG tail (cbhg_tail_kernel) is part of the conv/linear GEMM family bench.py reports under conv_gemm_tc_kernel
if 'cbhg_tail_kernel' in out and 'conv_gemm_tc_kernel' in out:
    tl = out['cbhg_tail_kernel']
    base = out['conv_gemm_tc_kernel']
    for k in ('launches_per_step', 'dram_bytes_per_step', 'us_serialised'):
        base[k] += tl[k]
    base['tail_launches_per_step'] = tl['launches_per_step']
json.dump(out, open(f'profiles/{tag}_traffic.json', 'w'), indent=1)
