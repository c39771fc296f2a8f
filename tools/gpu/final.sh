set -x
export B200DP_TIMEOUT_S=30
timeout -k 5 90 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; tail -2 gpurun_out/r2_smoke.log
timeout -k 5 200 python -m pytest tests -m gpu -q -p no:cacheprovider 2>&1 | tail -60 > gpurun_out/r2_pytest_n1_final.log
tail -3 gpurun_out/r2_pytest_n1_final.log
timeout -k 5 200 python bench.py --steps 30 --warmup 6 --kernel-detail gpurun_out/r2_kernel_detail_final.txt > gpurun_out/r2_bench_n1_final.json 2> gpurun_out/r2_bench_n1_final.err
python tools/bn_one.py --reps 2 64:112 256:14 512:7 > gpurun_out/plain_bn_one.log 2>&1 &&
timeout -k 5 150 ncu --set full --clock-control none --import-source on -k regex:"bn_stats|bn_bwd_reduce" -s 6 -c 6 -o gpurun_out/r2_prof_bn_reductions python tools/bn_one.py --reps 2 64:112 256:14 512:7 > gpurun_out/ncu_bn.log 2>&1
tail -3 gpurun_out/ncu_bn.log
