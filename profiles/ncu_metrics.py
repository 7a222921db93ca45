"""Print the headline metrics of an .ncu-rep: python profiles/ncu_metrics.py <file.ncu-rep>"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines())); hdr = rows[0]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct',
        'launch__registers_per_thread', 'gpu__dram_throughput.avg', 'sm__throughput.avg.pct',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'lts__t_sectors_srcunit_tex_op_write.sum', 'sm__cycles_elapsed.max',
        'lts__throughput.avg.pct', 'l1tex__throughput.avg.pct', 'lts__t_sector_hit_rate', 'lts__t_bytes.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum', 'lts__t_sectors.sum',
        'sm__inst_executed_pipe_lsu', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum']
for i, h in enumerate(hdr):
    if any(h.startswith(k) for k in keys):
        print(h, '|', rows[1][i], '|', [r[i] for r in rows[2:]])
