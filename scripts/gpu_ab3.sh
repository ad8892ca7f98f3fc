# A/B: early gather (ab/eg.so), 4 resident blocks for the MASK kernels (ab/mb4.so) vs the in-tree build
mkdir -p gpurun_out
{ echo "== spinodal, wave"; K=wave EXTRA="" bash scripts/gpu_ab.sh
  echo "== porous, wave"; K=wave EXTRA="--workload porous" bash scripts/gpu_ab.sh
  echo "== porous, two_pass"; K=two_pass EXTRA="--workload porous" bash scripts/gpu_ab.sh; } 2>&1 | tee gpurun_out/ab3.txt
DLBC_CUDA_LIB=$PWD/ab/eg.so timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/pytest_gpu_eg.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_eg.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu_eg.log | head -20
