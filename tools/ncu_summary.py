#!/usr/bin/env python
"""ncu report -> committed CSV summary (one block per captured kernel launch) + DRAM traffic per launch.
    python tools/ncu_summary.py gpurun_out/X.ncu-rep profiles/Y_ncu_summary.csv "how it was captured" """
import csv, json, subprocess, sys
KEEP = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit',
        'launch__shared_mem_per_block_dynamic', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed', 'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__throughput.avg.pct',
        'smsp__average_warps_issue_stalled', 'sass__inst_executed_local', 'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum',
        'smsp__sass_thread_inst_executed_op_dadd_pred_on.sum', 'smsp__sass_thread_inst_executed_op_dmul_pred_on.sum',
        'smsp__inst_executed_op_tma_ld.sum', 'sm__throughput.avg.pct', 'sm__icc_request_hit_rate.pct', 'lts__t_sectors_op_read.sum',
        'lts__t_sectors_op_write.sum']
rep, out, how = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
H, U = rows[0], rows[1]
traffic = {}
with open(out, 'w') as f:
    f.write("# %s\n# (captured under the profiler: durations are cold-cache / serialised -- bench.py's CUDA-event numbers are the timings of record)\n" % how)
    for V in rows[2:]:
        d = dict(zip(H, V))
        name = d.get('Kernel Name', '?')
        f.write("\n## %s\nmetric,unit,value\n" % name)
        for h, u, v in zip(H, U, V):
            if any(k in h for k in KEEP):
                f.write("%s,%s,%s\n" % (h, u, v))
        mult = {'Mbyte': 1e6, 'Gbyte': 1e9, 'Kbyte': 1e3, 'byte': 1, 'Tbyte': 1e12}
        rd = float(d['dram__bytes_read.sum'].replace(',', '')) * mult[U[H.index('dram__bytes_read.sum')]]
        wr = float(d['dram__bytes_write.sum'].replace(',', '')) * mult[U[H.index('dram__bytes_write.sum')]]
        traffic[name.split('(')[0].strip()] = dict(dram_bytes_read=rd, dram_bytes_write=wr, traffic_bytes_per_launch=rd + wr)
        print(name[:60], d['gpu__time_duration.sum'], 'issue', d.get('smsp__issue_active.avg.pct_of_peak_sustained_active'),
              'fp64', d.get('sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active'), 'MB', round((rd + wr) / 1e6, 1))
print(json.dumps(traffic))
