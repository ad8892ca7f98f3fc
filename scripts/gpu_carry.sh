# A/B of the solid-carry build (ab/carry.so, -DDLBC_SOLID_COPY=1) on the porous workload + the GPU suite with it
mkdir -p gpurun_out
EXTRA="--workload porous" bash scripts/gpu_ab.sh 2>&1 | tee gpurun_out/ab_carry.txt
DLBC_CUDA_LIB=$PWD/ab/carry.so timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/pytest_gpu_carry.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_carry.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu_carry.log | head -20
grep -q "rc=0" gpurun_out/pytest_gpu_carry.log || tail -40 gpurun_out/pytest_gpu_carry.log
