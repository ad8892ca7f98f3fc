# A/B: psi halo rows shared between neighbouring bands (ab/share.so) vs the in-tree build; GPU suite with it first
mkdir -p gpurun_out
DLBC_CUDA_LIB=$PWD/ab/share.so timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/pytest_gpu_share.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_share.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu_share.log | head -20
grep -q "rc=0" gpurun_out/pytest_gpu_share.log || { tail -30 gpurun_out/pytest_gpu_share.log; exit 1; }
{ echo "== spinodal, wave"; K=wave EXTRA="" bash scripts/gpu_ab.sh
  echo "== porous, wave"; K=wave EXTRA="--workload porous" bash scripts/gpu_ab.sh; } 2>&1 | tee gpurun_out/ab_share_halo.txt
