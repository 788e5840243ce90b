"""Print the key metrics of every kernel in an .ncu-rep (run here, no GPU needed): python tools/ncu_summary.py rep [extra-metric-substrings]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__cluster_size',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__cycles_elapsed.avg', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'smsp__inst_executed.sum',
        'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_alu.sum',
        'l1tex__data_pipe_lsu_wavefronts.sum', 'smsp__cycles_active.avg']
extra = sys.argv[2:]
for r in rows[2:]:
    if '-nan' in r or 'nan' in r[h.index('sm__cycles_elapsed.avg')]:
        continue
    print('-----', r[h.index('Kernel Name')][:100])
    for w in want:
        if w in h:
            i = h.index(w); print(f"  {w:70s} {r[i]:>22s} {units[i]}")
    for i, n in enumerate(h):
        if 'issue_stalled' in n and 'pcsamp' in n and 'not_issued' not in n and r[i] not in ('0', 'nan', '-nan'):
            print(f"  {n:70s} {r[i]:>22s}")
        elif extra and any(x in n for x in extra):
            print(f"  {n:70s} {r[i]:>22s} {units[i]}")
