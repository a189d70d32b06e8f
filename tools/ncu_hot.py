"""Top SASS lines by warp-stall samples, with the dominant stall reason and the executed-instruction count:
python tools/ncu_hot.py rep [kernel-regex] [N]"""
import csv, subprocess, sys
rep = sys.argv[1]; pat = sys.argv[2] if len(sys.argv) > 2 else ''; N = int(sys.argv[3]) if len(sys.argv) > 3 else 25
cmd = ['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass']
if pat: cmd += ['--kernel-name', 'regex:' + pat]
out = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == 'Kernel Name':
        name = rows[i][1]; h = rows[i + 1]; j = i + 2
        data = []
        while j < len(rows) and not (rows[j] and rows[j][0] == 'Kernel Name'):
            data.append(rows[j]); j += 1
        col = {c: k for k, c in enumerate(h)}
        stalls = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
        tot = sum(float(r[col['# Samples']]) for r in data if len(r) > col['# Samples']) or 1
        print(name[:100], 'samples', tot)
        agg = {c: sum(float(r[col[c]]) for r in data if len(r) > col[c]) for c in stalls}
        print('  stall mix:', ', '.join('%s %.1f%%' % (c[6:], 100 * v / tot) for c, v in sorted(agg.items(), key=lambda x: -x[1])[:7]))
        for r in sorted(data, key=lambda r: -float(r[col['# Samples']]))[:N]:
            s = float(r[col['# Samples']])
            top = max(stalls, key=lambda c: float(r[col[c]]))
            print('  %5.2f%%  %-10s exec %8s  %s' % (100 * s / tot, top[6:], r[col['Instructions Executed']], r[col['Source']].strip()[:110]))
        i = j
    else:
        i += 1
