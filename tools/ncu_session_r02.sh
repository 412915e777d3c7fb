#!/bin/bash
# Round-2 profiling session (run under gpurun on ONE B200): launch list of the bench command, a full capture of the two
# blind-rotation kernels the verdict names (2_2 tensor-memory kernel, N = 8192 kernel), and the FP64-pipe counters of the
# DFMA / DMMA probe.  Outputs under gpurun_out/; the summaries are distilled into profiles/ by tools/ncu_brief.py.
set -x
cd "$(dirname "$0")/.."
CMD="python bench.py --steps 2 --warmup 3 --extras none --no-cpu-baseline"
$CMD > gpurun_out/r02_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r02_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:blind_rotate_tmem -s 3 -c 1 -o gpurun_out/r02_prof_tmem -f \
    --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum \
    $CMD > gpurun_out/r02_ncu_tmem.log 2>&1
CMD3="python bench.py --steps 2 --warmup 3 --extras shortint_3_3 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:blind_rotate_big -s 2 -c 1 -o gpurun_out/r02_prof_big -f $CMD3 > gpurun_out/r02_ncu_big.log 2>&1
ncu --clock-control none --metrics sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_fp64.sum,sm__inst_executed_pipe_tensor.sum,smsp__inst_executed.sum,gpu__time_duration.sum \
    --csv --log-file gpurun_out/r02_pipe_probe_ncu.csv ./tools/micro/pipe_probe > gpurun_out/r02_pipe_probe_under_ncu.log 2>&1
ls -la gpurun_out/
