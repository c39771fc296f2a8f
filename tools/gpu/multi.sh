# usage: bash tools/gpu/multi.sh N   (inside gpurun --gpus N)
N=$1
export B200DP_TIMEOUT_S=30
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
nvidia-smi topo -m > gpurun_out/r2_topo_n$N.txt 2>&1
timeout -k 5 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -s -p no:cacheprovider > gpurun_out/r2_multi_n$N.log 2>&1
tail -5 gpurun_out/r2_multi_n$N.log
timeout -k 5 300 $TR tools/allreduce_sweep.py > gpurun_out/r2_allreduce_sweep_n$N.jsonl 2> gpurun_out/r2_allreduce_sweep_n$N.err
SWEEP_NVLS=1 timeout -k 5 300 $TR tools/allreduce_sweep.py > gpurun_out/r2_allreduce_sweep_nvls_n$N.jsonl 2> gpurun_out/r2_allreduce_sweep_nvls_n$N.err
tail -3 gpurun_out/r2_allreduce_sweep_nvls_n$N.err
timeout -k 5 300 $TR tools/stat_exchange_sweep.py > gpurun_out/r2_stat_exchange_n$N.jsonl 2> gpurun_out/r2_stat_exchange_n$N.err
timeout -k 5 400 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
tail -c 400 gpurun_out/r2_bench_n$N.err
