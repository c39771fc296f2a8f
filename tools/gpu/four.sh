# usage: bash tools/gpu/eight.sh N    (inside gpurun --gpus N): world-N tests, config-5 sweeps, bench lines
N=$1
export B200DP_TIMEOUT_S=30
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
nvidia-smi topo -m > gpurun_out/r2_topo_n$N.txt 2>&1
timeout -k 5 420 python -m pytest tests/test_gpu_multi.py -m gpu -v -s -p no:cacheprovider -k "$N" > gpurun_out/r2_multi_n$N.log 2>&1
grep -v "frame #\|should dump\|^$\|Exception raised from\|recvValue\|sendBytes\|c10::Error\|unknown function\|UserWarning\|Consider using\|key = (\|moved = float" gpurun_out/r2_multi_n$N.log | tail -25
timeout -k 5 200 $TR tools/allreduce_sweep.py > gpurun_out/r2_allreduce_sweep_n$N.jsonl 2> gpurun_out/r2_allreduce_sweep_n$N.err
SWEEP_NVLS=1 timeout -k 5 200 $TR tools/allreduce_sweep.py > gpurun_out/r2_allreduce_sweep_nvls_n$N.jsonl 2> gpurun_out/r2_allreduce_sweep_nvls_n$N.err
timeout -k 5 150 $TR tools/stat_exchange_sweep.py > gpurun_out/r2_stat_exchange_n$N.jsonl 2> gpurun_out/r2_stat_exchange_n$N.err
timeout -k 5 300 $TR bench.py --gpus $N --steps 25 --warmup 6 --no-cpu-baseline --kernel-detail gpurun_out/r2_kernel_detail_n$N.txt > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
tail -c 300 gpurun_out/r2_bench_n$N.err
