mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o /tmp/ubench_tma scripts/ubench_tma.cu || exit 1
UB_MODE0=2 UB_QUICK=1 UB_SHIFT=0 timeout 120 compute-sanitizer /tmp/ubench_tma 2>&1 | head -40 > gpurun_out/ubench_sanitize.txt
cat gpurun_out/ubench_sanitize.txt
echo "--- aligned"; UB_MODE0=2 UB_SHIFT=0 timeout 200 /tmp/ubench_tma 2>&1 | tee gpurun_out/ubench_tma_tensor_aligned.txt | head -60
echo "--- shifted"; UB_MODE0=2 UB_SHIFT=1 timeout 200 /tmp/ubench_tma 2>&1 | tee gpurun_out/ubench_tma_tensor_shift.txt | head -60
