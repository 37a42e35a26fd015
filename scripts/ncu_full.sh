#!/bin/bash
# ncu --set full of the step's kernels (one launch each, steady state), after a plain run.  KERN: regex of kernel names.
set -u
mkdir -p gpurun_out
TAG=${TAG:-full}
CMD="python scripts/kernel_times.py ${T:-240} ${N:-10000} bf16x3"
timeout 300 $CMD > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${KERN:-step_reduce} -s ${SKIP:-3} -c ${COUNT:-1} \
   -o gpurun_out/${TAG}_prof -f $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/${TAG}_ncu.log
ncu -i gpurun_out/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
python - <<PY
import csv
rows = list(csv.reader(open("gpurun_out/${TAG}_raw.csv")))
hdr, units, vals = rows[0], rows[1], rows[2:]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_fp64.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_lg.sum", "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
        "lts__t_sectors_srcunit_tex_op_read.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"]
for v in vals:
    name = v[hdr.index("Kernel Name")][:50]
    print("==", name)
    for i, h in enumerate(hdr):
        if h in want or h.startswith("smsp__average_warps_issue_stalled") or h.startswith("smsp__average_warp_latency_issue_stalled"):
            try:
                if h.startswith("smsp__average") and float(v[i].replace(",", "")) < 0.3: continue
            except ValueError: pass
            print(f"  {h} = {v[i]} {units[i]}")
PY
