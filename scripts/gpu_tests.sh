# full -m gpu suite, fail fast, bounded
mkdir -p gpurun_out
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout ${T:-600} python -m pytest tests -m gpu -x -q --timeout 120 $PYTEST_ARGS > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu.log | head -20
grep -q "rc=0" gpurun_out/pytest_gpu.log || tail -40 gpurun_out/pytest_gpu.log
