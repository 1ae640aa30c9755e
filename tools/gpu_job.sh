cd /root/repo
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu_final_1gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r02_pytest_gpu_final_1gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke_final.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err; echo "bench rc=$?"; tail -1 gpurun_out/r02_bench_final.json | cut -c1-260
timeout 600 python bench.py --workload ac --steps 10 --warmup 3 > gpurun_out/r02_bench_ac_final.json 2> gpurun_out/r02_bench_ac_final.err; echo "bench ac rc=$?"; tail -1 gpurun_out/r02_bench_ac_final.json | cut -c1-260
