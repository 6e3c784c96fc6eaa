#!/bin/bash
# Runs the measurements a round commits under profiles/: bench lines, ncu launch lists and ncu --set full
# captures of the two dominant kernels. Meant for one GPU box:  gpurun -- 'bash tools/round_profile.sh r1e'
tag=${1:-rX}
out=gpurun_out
mkdir -p $out
python bench.py --steps 20 --warmup 3 > $out/${tag}_bench_c2.json 2> $out/${tag}_bench_c2.err
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference_c2.json 2> $out/${tag}_bench_reference_c2.err
for w in c1 c3 c5; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu > $out/${tag}_bench_$w.json 2> $out/${tag}_bench_$w.err
done
for w in c2 c3 c5; do
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches_$w.csv \
      python bench.py --workload $w --steps 2 --warmup 3 --no-cpu > $out/${tag}_ncu_$w.log 2>&1
  python tools/launch_summary.py $out/${tag}_launches_$w.csv > $out/${tag}_launches_$w.txt 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:wfShadeKernel -s 12 -c 1 -f -o $out/${tag}_wfshade_c2 \
    python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu > $out/${tag}_ncu_shade.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:wfTrace -s 12 -c 1 -f -o $out/${tag}_wftrace_c2 \
    python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu > $out/${tag}_ncu_trace.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:wfTrace -s 13 -c 1 -f -o $out/${tag}_wftrace_shadow_c3 \
    python bench.py --workload c3 --steps 1 --warmup 3 --no-cpu > $out/${tag}_ncu_trace3.log 2>&1
METRICS="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,launch__occupancy_limit_registers,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.max,dram__throughput.avg.pct_of_peak_sustained_elapsed,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"
for r in wfshade_c2 wftrace_c2 wftrace_shadow_c3; do
  if [ -f $out/${tag}_$r.ncu-rep ]; then
    ncu -i $out/${tag}_$r.ncu-rep --page raw --metrics $METRICS > $out/${tag}_$r.txt 2>&1
  fi
done
python tools/ncu_lines.py $out/${tag}_wfshade_c2.ncu-rep wfShadeKernelILb1ELb0 --top 45 > $out/${tag}_wfshade_c2_lines.txt 2>&1
python tools/ncu_lines.py $out/${tag}_wftrace_c2.ncu-rep wfTraceSimpleKernel --top 30 > $out/${tag}_wftrace_c2_lines.txt 2>&1
ls -la $out | tail -40
