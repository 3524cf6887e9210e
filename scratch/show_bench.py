"""Print the headline and the per-launch table of a bench.py JSON line: python scratch/show_bench.py FILE"""
import json
import sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches', 'kernel_ms_sum')}, 'e2e', d['e2e']['value'])
r = d['roofline']
print('gemm ms', r['ms_per_step'], 'TF', round(r['achieved'], 1), 'frac', round(r['frac'], 3), 'f16', r.get('f16_flop_share'))
for k in d['kernels']:
    print(f"{k.get('side', ''):6s} {k.get('label', '')[:70]:70s} {k.get('kind', ''):5s} {k.get('ms')} {k.get('tflops', '')} {k.get('f16_share', '')}")
