#!/usr/bin/env python
"""Turn gpurun_out/prof_decode_<wl>.ncu-rep + launches_<wl>.csv into the committed summaries under profiles/."""
import collections, csv, json, subprocess
keep = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit', 'launch__shared_mem_per_block_dynamic', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__average_warps_issue_stalled', 'sass__inst_executed_local',
        'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum', 'smsp__sass_thread_inst_executed_op_dadd_pred_on.sum',
        'smsp__sass_thread_inst_executed_op_dmul_pred_on.sum', 'smsp__inst_executed_op_tma_ld.sum', 'sm__throughput.avg.pct', 'sm__icc_request_hit_rate.pct']
desc = {'target': 'k_decode<12,HSCN,binomial,reference,RAW (columns read in place)> on 10,000 cells x 5,700 bins (57 M cell-bins)',
        'cfg3': 'k_decode<12,ASCN,beta-binomial,reference,LDG> on 10,000 cells x 5,700 bins (57 M cell-bins)'}
traffic = {}
for wl in ('target', 'cfg3'):
    raw = subprocess.run(['ncu', '-i', f'gpurun_out/prof_decode_{wl}.ncu-rep', '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    H, U, V = rows[0], rows[1], rows[2]
    out = f'profiles/r01_k_decode_{wl}_ncu_summary.csv'
    d = {}
    with open(out, 'w') as f:
        f.write(f"# ncu --set full --clock-control none --import-source on -k regex:k_decode -s 3 -c 1 ; {desc[wl]}\n"
                "# (captured under the profiler: durations are cold-cache/serialised -- bench.py's CUDA-event numbers are the timings of record)\nmetric,unit,value\n")
        for h, u, v in zip(H, U, V):
            d[h] = v
            if any(k in h for k in keep):
                f.write(f"{h},{u},{v}\n")
    unit = U[H.index('dram__bytes_read.sum')]
    mult = {'Mbyte': 1e6, 'Gbyte': 1e9, 'Kbyte': 1e3, 'byte': 1}[unit]
    rd = float(d['dram__bytes_read.sum'].replace(',', '')) * mult
    wr = float(d['dram__bytes_write.sum'].replace(',', '')) * mult
    traffic[wl] = {'dram_bytes_read': rd, 'dram_bytes_write': wr, 'traffic_bytes_per_launch': rd + wr, 'kernel': 'k_decode', 'source': out}
    print(wl, d['gpu__time_duration.sum'], 'fp64', d['sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active'], 'issue',
          d['smsp__issue_active.avg.pct_of_peak_sustained_active'], 'smem', d['l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed'],
          'regs', d['launch__registers_per_thread'], 'inst', d['smsp__inst_executed.sum'], 'MB', (rd + wr) / 1e6)
    rows = list(csv.reader(open(f'gpurun_out/launches_{wl}.csv')))
    hdr = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
    Hh = rows[hdr]; ki = Hh.index('Kernel Name'); vi = Hh.index('Metric Value')
    agg = collections.defaultdict(list)
    for r in rows[hdr + 1:]:
        if len(r) > vi:
            agg[r[ki][:70]].append(float(r[vi].replace(',', '')))
    tot = sum(sum(v) for v in agg.values())
    with open(f'profiles/r01_launches_{wl}.txt', 'w') as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none -c 400 python bench.py --steps 3 --warmup 3 --no-extras --workload {wl}\n"
                "# per-kernel device time over the whole process (warm-up + timed steps + e2e chunk launches); shares, not absolutes\n")
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"{k:70s} n={len(v):4d} total_ms={sum(v)/1e6:9.3f} mean_us={sum(v)/len(v)/1e3:9.1f} share={sum(v)/tot*100:5.1f}%\n")
json.dump(traffic, open('profiles/r01_k_decode_traffic.json', 'w'), indent=1)
