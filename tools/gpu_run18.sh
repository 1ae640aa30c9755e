# 1 GPU: full GPU suite, default bench (with CPU baseline), launch list, env-kernel capture with FP32 pipe metrics
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r01_pytest_gpu_v8.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01_pytest_gpu_v8.log
timeout 900 python bench.py > gpurun_out/r01_bench_v8.json 2> gpurun_out/r01_bench_v8.err; echo "bench rc=$?"; tail -2 gpurun_out/r01_bench_v8.err
grep -h '^{' gpurun_out/r01_bench_v8.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value']), d['phase_ms'], d['roofline_fp32'], d['e2e']['value'], d['cpu_baseline'])"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r01_launches_v8.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_v8.log 2>&1
tail -2 gpurun_out/ncu_v8.log | cut -c1-200
M=smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__cycles_active.avg
timeout 900 ncu --set full --metrics $M --clock-control none --import-source on -k regex:env_step_kernel --launch-skip 110 -c 1 -o gpurun_out/r01_envstep_v8 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_envstep_v8.log 2>&1
tail -3 gpurun_out/ncu_envstep_v8.log | cut -c1-200
