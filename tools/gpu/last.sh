export B200DP_TIMEOUT_S=30
timeout -k 5 60 python -m pytest tests/test_gpu_ddp.py -m gpu -q -p no:cacheprovider -k "pickle" 2>&1 | tail -3 > gpurun_out/r2_pytest_pickle.log; cat gpurun_out/r2_pytest_pickle.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-kernel-pass > gpurun_out/plain_launches.log 2>&1 &&
timeout -k 5 110 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r2_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-kernel-pass > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log; wc -l gpurun_out/r2_launches_bench_steps2.csv
