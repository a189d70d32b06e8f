"""Print the handful of ncu metrics we track, per profiled launch, from an .ncu-rep (raw page csv)."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'sm__inst_executed_pipe_tensor', 'sm__pipe_tensor', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__average_warps_issue_stalled',
        'sm__throughput.avg.pct', 'l1tex__throughput.avg.pct', 'lts__throughput.avg.pct', 'sm__cycles_elapsed.avg',
        'smsp__cycles_active.avg', 'lts__t_sectors_srcunit_tex_op_read.sum', 'lts__t_sector_hit_rate',
        'l1tex__lsu_writeback', 'sm__memory_throughput', 'l1tex__data_pipe']
for r in rows[2:]:
    print('=' * 100)
    for i, n in enumerate(h):
        if any(n == k or n.startswith(k) for k in keys):
            if r[i] not in ('', '0', 'n/a'):
                print('%-90s %-10s %s' % (n, u[i], r[i]))
