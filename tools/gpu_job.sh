cd /root/repo
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu_final_1gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_final_1gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke_final.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err; echo "bench rc=$?"; tail -1 gpurun_out/r02_bench_final.json | cut -c1-260
timeout 600 python bench.py --workload ac --steps 10 --warmup 3 > gpurun_out/r02_bench_ac_final.json 2> gpurun_out/r02_bench_ac_final.err; echo "bench ac rc=$?"; tail -1 gpurun_out/r02_bench_ac_final.json | cut -c1-260
timeout 600 python bench.py --impl reference --workload ac --steps 1 --warmup 1 > gpurun_out/r02_bench_ac_ref_final.json 2> gpurun_out/r02_bench_ac_ref_final.err; echo "ref ac rc=$?"; tail -1 gpurun_out/r02_bench_ac_ref_final.json | cut -c1-300
timeout 600 python tools/bench_configs.py 5 3 C2,C3,HC > gpurun_out/r02_bench_configs.jsonl 2> gpurun_out/r02_bench_configs.err; echo "configs rc=$?"; cut -c1-330 gpurun_out/r02_bench_configs.jsonl
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_launches_ac_final.csv python bench.py --workload ac --steps 2 --warmup 2 --no-cpu-baseline --profile-iteration > gpurun_out/ncu_launches_ac.log 2>&1
wc -l gpurun_out/r02_launches_ac_final.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:env_step_kernel --launch-skip 100 -c 1 -f -o gpurun_out/r02_envstep_halfcheetah python tools/profile_ac.py 8192 1 > gpurun_out/ncu_ac_env.log 2>&1; tail -1 gpurun_out/ncu_ac_env.log
