set -x
export B200DP_TIMEOUT_S=30
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout -k 10 1200 python -m pytest tests -m gpu -q -p no:cacheprovider 2>&1 | tail -80 > gpurun_out/r2_pytest_n1_a.log
tail -5 gpurun_out/r2_pytest_n1_a.log
timeout 300 python tools/bn_bench.py --warm --one-kernel-mb 0 0 > gpurun_out/r2_bn_warm_two_kernels.txt 2>&1
timeout 300 python tools/bn_bench.py --warm --one-kernel-mb 56 40 > gpurun_out/r2_bn_warm_one_kernel_56_40.txt 2>&1
timeout 300 python tools/bn_bench.py --warm --one-kernel-mb 110 110 --cluster 8 > gpurun_out/r2_bn_warm_one_kernel_110_cs8.txt 2>&1
timeout 300 python tools/bn_bench.py --one-kernel-mb 56 40 > gpurun_out/r2_bn_cold_one_kernel_56_40.txt 2>&1
timeout 600 python bench.py --steps 30 --warmup 5 --kernel-detail gpurun_out/r2_kernel_detail_a.txt > gpurun_out/r2_bench_n1_a.json 2> gpurun_out/r2_bench_n1_a.err
tail -c 600 gpurun_out/r2_bench_n1_a.err
B200DP_BN_FUSED=0 timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-kernel-pass > gpurun_out/r2_bench_n1_a_two_kernels.json 2>&1
cat gpurun_out/r2_bench_n1_a_two_kernels.json | tail -c 700
