# usage: N=<gpus> bash scripts/gpu_multi_tests_only.sh   (under gpurun --gpus N): decomposition tests only, log kept
mkdir -p gpurun_out
N=${N:-2}
timeout 400 python -m pytest tests/test_multi_gpu.py -q --timeout 200 -rs > gpurun_out/pytest_multi_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi_n$N.log
tail -5 gpurun_out/pytest_multi_n$N.log
