export B200DP_TIMEOUT_S=30
timeout -k 5 420 python -m pytest tests/test_gpu_multi.py -m gpu -v -s -p no:cacheprovider -k "8" > gpurun_out/r2_multi_n8.log 2>&1
timeout -k 5 120 python -m pytest tests/test_gpu_multi.py -m gpu -v -s -p no:cacheprovider -k "4-sgd" > gpurun_out/r2_multi_n4_sgd.log 2>&1
grep -v "frame #\|should dump\|^$\|Exception raised from\|recvValue\|sendBytes\|c10::Error\|unknown function\|UserWarning\|Consider using\|key = (\|moved = float" gpurun_out/r2_multi_n8.log | tail -25
tail -3 gpurun_out/r2_multi_n4_sgd.log
