mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_d2q9.py tests/test_params_replay.py -m gpu -q --timeout 120 > gpurun_out/pytest_d2q9.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_d2q9.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_d2q9.log | head -30
grep -E "^E  " gpurun_out/pytest_d2q9.log | head -30
