set -x
export B200DP_TIMEOUT_S=30
timeout -k 5 400 python -m pytest tests -m gpu -q -p no:cacheprovider 2>&1 | tail -300 > gpurun_out/r2_pytest_n1_c.log
tail -3 gpurun_out/r2_pytest_n1_c.log
timeout 300 python tools/bn_bench.py --warm --chain 8 > gpurun_out/r2_bn_warm8_default.txt 2>&1
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --kernel-detail gpurun_out/r2_kernel_detail_c.txt > gpurun_out/r2_bench_n1_c.json 2> gpurun_out/r2_bench_n1_c.err
