mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_params_replay.py -m gpu -q --timeout 60 -k "init_2d" > gpurun_out/pytest_init2d.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_init2d.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_init2d.log | head -40
grep -E "^E  " gpurun_out/pytest_init2d.log | head -10
