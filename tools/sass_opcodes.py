"""Per-kernel SASS opcode counts of libsmplb200.so (what proves tcgen05 / TMEM / bulk-copy code): markdown table.
usage: python tools/sass_opcodes.py > profiles/r02_sass_opcodes.md"""
import collections, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, '3d-pose-estimation_b200', 'smplb200', 'libsmplb200.so')
out = subprocess.run(['cuobjdump', '-sass', so], capture_output=True, text=True).stdout
ops = ['UTCHMMA', 'UTCHMMA.2CTA', 'LDTM', 'STTM', 'UBLKCP', 'UTCBAR', 'SYNCS', 'LDGSTS', 'HMMA', 'F2FP.BF16', 'F2FP.F16', 'STG', 'LDG', 'STS', 'LDS']
cur, cnt = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r'Function : (\S+)', line)
    if m:
        cur = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip().split('(')[0]
        cnt[cur] = collections.Counter()
        continue
    m = re.match(r'\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', line)
    if m and cur:
        op = m.group(1)
        for o in ops:
            if op == o or op.startswith(o + '.'):
                cnt[cur][o] += 1
head = subprocess.run(['git', '-C', ROOT, 'rev-parse', '--short', 'HEAD'], capture_output=True, text=True).stdout.strip()
print('# SASS opcode counts per kernel (`cuobjdump -sass libsmplb200.so`, built from %s + working tree, sm_100a)\n' % head)
print('UTCHMMA = tcgen05.mma (kind::f16), LDTM/STTM = tcgen05.ld/st, UBLKCP = cp.async.bulk (TMA engine), UTCBAR = tcgen05.commit,')
print('SYNCS = mbarrier ops, LDGSTS = cp.async, HMMA = legacy mma.sync (none expected).\n')
print('| kernel | ' + ' | '.join(ops) + ' |')
print('|---|' + '---|' * len(ops))
for k, c in cnt.items():
    if any(c[o] for o in ops[:9]) or 'kernel' in k:
        print('| `%s` | ' % k.replace('smplb::', '') + ' | '.join(str(c[o]) for o in ops) + ' |')
