python tools/bn_one.py --reps 1 64:112 > gpurun_out/plain_bn_large.log 2>&1 &&
timeout -k 3 40 ncu --set full --clock-control none -k regex:"bn_stats|bn_bwd_reduce" -c 2 -o gpurun_out/r2_prof_bn_reductions_large python tools/bn_one.py --reps 1 64:112 > gpurun_out/ncu_bn_large.log 2>&1
tail -2 gpurun_out/ncu_bn_large.log
