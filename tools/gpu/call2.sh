set -x
export B200DP_TIMEOUT_S=30
timeout -k 10 900 python -m pytest tests -m gpu -q -p no:cacheprovider 2>&1 | tail -400 > gpurun_out/r2_pytest_n1_b.log
tail -5 gpurun_out/r2_pytest_n1_b.log
