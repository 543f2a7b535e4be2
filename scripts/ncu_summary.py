"""Key metrics + stall breakdown of one kernel from an .ncu-rep (read here, no GPU).
   python scripts/ncu_summary.py gpurun_out/prof_rti.ncu-rep [out.csv]"""
import csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, val = rows[0], rows[1], rows[2]
d = {h: (u, v) for h, u, v in zip(hdr, units, val)}
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size',
        'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic',
        'launch__occupancy_limit_registers', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tc.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tma.avg.pct_of_peak_sustained_active',
        'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed',
        'smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed',
        'smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'dram__bytes_read.sum',
        'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__average_warp_latency_per_inst_issued.ratio', 'sass__inst_executed_local_loads',
        'sass__inst_executed_local_stores']
stalls = sorted(((float(v[1]), h) for h, v in d.items()
                 if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio')),
                reverse=True)
out = [('metric', 'unit', 'value')] + [(k, *d.get(k, ('', 'n/a'))) for k in keys]
out += [(h.replace('smsp__average_warps_issue_stalled_', 'stall_').replace('_per_issue_active.ratio', ''),
         'warps/issue', f'{v:.3f}') for v, h in stalls if v > 0.004]
for r in out:
    print(f'{r[0][:78]:78s} {r[1]:16s} {r[2]}')
if len(sys.argv) > 2:
    csv.writer(open(sys.argv[2], 'w')).writerows(out)
