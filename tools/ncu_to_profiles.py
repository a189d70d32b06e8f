"""Turn an `ncu --set full` report of the forward step into the two small records bench.py reads, stamped with the git
blob hashes of the kernel sources they were captured from (bench.py prints null when a source has changed since):
    profiles/skin_traffic.json       dram bytes per launch of skin_tc_kernel<0>        -> roofline.traffic
    profiles/blend_tensor_pipe.json  tensor-pipe activity of blend_tc_kernel           -> roofline_gemm.tensor_pipe_active
usage: python tools/ncu_to_profiles.py gpurun_out/<rep>.ncu-rep <batch> [summary.md]"""
import csv
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, '3d-pose-estimation_b200', 'csrc')


def blob_sha(path):
    data = open(path, 'rb').read()
    return hashlib.sha1(b'blob %d\0' % len(data) + data).hexdigest()


def sources(names):
    return {n: blob_sha(os.path.join(CSRC, n)) for n in names}


def main():
    rep, batch = sys.argv[1], int(sys.argv[2])
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    col = {n: i for i, n in enumerate(rows[0])}
    unit = {n: rows[1][i] for n, i in col.items()}
    head = subprocess.run(['git', '-C', ROOT, 'rev-parse', '--short', 'HEAD'], capture_output=True, text=True).stdout.strip()
    scale = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0, 'Tbyte': 1e12}
    tscale = {'ms': 1.0, 'us': 1e-3, 'ns': 1e-6, 's': 1e3}

    def f(r, name):
        return float(r[col[name]].replace(',', ''))

    md = []
    for r in rows[2:]:
        name = r[col['Kernel Name']]
        rd = f(r, 'dram__bytes_read.sum') * scale[unit['dram__bytes_read.sum']]
        wr = f(r, 'dram__bytes_write.sum') * scale[unit['dram__bytes_write.sum']]
        dur = f(r, 'gpu__time_duration.sum') * tscale[unit['gpu__time_duration.sum']]
        tens = f(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed')
        ghz = f(r, 'sm__cycles_elapsed.avg.per_second')
        md.append('| `%s` | %.3f | %.3f | %.3f | %.1f | %.2f | %s |' % (
            name.split('(')[0], dur, rd / 1e9, wr / 1e9, tens, ghz, r[col['launch__registers_per_thread']]))
        if 'skin_tc_kernel<0>' in name:
            rec = {'kernel': 'skin_tc_kernel<0>', 'batch': batch, 'dram_bytes_per_launch': int(rd + wr),
                   'dram_bytes_read': int(rd), 'dram_bytes_write': int(wr), 'duration_ms_under_ncu': dur,
                   'report': os.path.basename(rep), 'captured_at_commit': head,
                   'sources': sources(['skin_tc.cu', 'tc_common.cuh', 'common.cuh'])}
            json.dump(rec, open(os.path.join(ROOT, 'profiles', 'skin_traffic.json'), 'w'), indent=1)
        if 'blend_tc_kernel' in name:
            rec = {'kernel': 'blend_tc_kernel', 'batch': batch, 'tensor_pipe_active_pct': tens, 'sm_clock_ghz': ghz,
                   'duration_ms_under_ncu': dur, 'dram_bytes_per_launch': int(rd + wr),
                   'report': os.path.basename(rep), 'captured_at_commit': head,
                   'sources': sources(['blend_tc.cu', 'tc_common.cuh', 'common.cuh'])}
            json.dump(rec, open(os.path.join(ROOT, 'profiles', 'blend_tensor_pipe.json'), 'w'), indent=1)
    table = ('| kernel | ms (under ncu) | DRAM read GB | DRAM write GB | tensor pipe active % | SM clock GHz | regs |\n'
             '|---|---|---|---|---|---|---|\n' + '\n'.join(md))
    print(table)
    if len(sys.argv) > 3:
        open(sys.argv[3], 'a').write('\n' + table + '\n')


if __name__ == '__main__':
    main()
