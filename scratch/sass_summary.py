"""Which tensor-core / TMA / cluster instructions each kernel of libftb200.so really contains (cuobjdump -sass).
    python scratch/sass_summary.py > profiles/r2_sass_summary.txt"""
import collections, re, subprocess
LIB = 'forwardtacotron_b200/libftb200.so'
PAT = re.compile(r'\b(UTCHMMA|UTCQMMA|UTCBAR|LDTM|STTM|UTMALDG|UTMASTG|HMMA|SYNCS|UCGABAR|STAS|LDGSTS|MUFU)([.\w]*)')
sass = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True).stdout
cnt = collections.OrderedDict()
fn = None
for line in sass.splitlines():
    m = re.search(r'Function : (\S+)', line)
    if m:
        name = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip()
        fn = re.sub(r'\(.*', '', name).replace('void ', '')
        cnt.setdefault(fn, collections.Counter())
        continue
    if fn is None:
        continue
    for m in PAT.finditer(line):
        key = m.group(1)
        suf = m.group(2)
        if key == 'HMMA':
            key += '.' + suf.split('.')[1] if '.' in suf else ''
        if '2CTA' in suf:
            key += '.2CTA'
        if 'MULTICAST' in suf:
            key += '.MULTICAST'
        if key == 'MUFU':
            key += suf.split('.')[0] if suf.startswith('.') and False else ('.' + suf.split('.')[1] if suf.count('.') else '')
        cnt[fn][key] += 1
print(f'# cuobjdump -sass {LIB}: instruction counts per kernel')
print('# UTCHMMA = tcgen05.mma (kind::f16 / tf32), LDTM = tcgen05.ld, UTMALDG = cp.async.bulk.tensor (TMA load), UTCBAR = tcgen05.commit,')
print('# HMMA.16816 = legacy mma.sync m16n8k16, SYNCS = mbarrier ops, UCGABAR = cluster barrier, STAS = st.async (DSMEM), MUFU.* = SFU')
for fn, c in cnt.items():
    if c:
        print(f'{fn:42s} ' + '  '.join(f'{k}={v}' for k, v in sorted(c.items())))
    else:
        print(f'{fn:42s} (none of the above)')
