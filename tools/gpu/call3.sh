export B200DP_TIMEOUT_S=30
for dt in f32 bf16; do for mode in 0 1 2; do
  B200DP_BN_FUSED=0 timeout 120 python tools/bn_steps.py $dt $mode
  B200DP_BN_FUSED=0 B200DP_BN_PDL_FIRST=0 timeout 120 python tools/bn_steps.py $dt $mode 2>&1 | sed 's/^/nopdl /' | grep -v " ok$"
  timeout 120 python tools/bn_steps.py $dt $mode 2>&1 | sed 's/^/fused /' | grep "1 call"
done; done > gpurun_out/r2_bn_steps.log 2>&1
