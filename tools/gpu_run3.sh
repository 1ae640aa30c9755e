set -x
PB_SPLITS=32 python tools/profile_update.py 1 32 > gpurun_out/plain.log 2>&1 && PB_SPLITS=32 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"fwd_bwd|wgrad|grad_reduce|adam2|epoch_gather|adv_stats" -s 140 -c 80 --csv --log-file gpurun_out/r01_update_launches_v2.csv python tools/profile_update.py 1 32 > gpurun_out/ncu.log 2>&1
cat gpurun_out/plain.log; tail -3 gpurun_out/ncu.log
