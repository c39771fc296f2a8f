N=$1
export B200DP_TIMEOUT_S=30
timeout -k 5 ${2:-700} python -m pytest tests/test_gpu_multi.py -m gpu -v -s -p no:cacheprovider > gpurun_out/r2_multi_n$N.log 2>&1
grep -v "frame #\|should dump\|^$\|Exception raised from\|recvValue\|sendBytes\|c10::Error\|unknown function\|UserWarning\|Consider using\|key = (" gpurun_out/r2_multi_n$N.log | tail -60
