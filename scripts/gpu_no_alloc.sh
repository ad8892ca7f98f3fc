# the allocation-free entry point: its test alone in a fresh process, then the whole -m gpu suite (it is collected last)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q --timeout 120 > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/pytest_no_alloc_alone.log
tail -4 gpurun_out/pytest_no_alloc_alone.log
timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu.log | head -20
grep -q "rc=0" gpurun_out/pytest_gpu.log || tail -40 gpurun_out/pytest_gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
