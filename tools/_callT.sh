mkdir -p gpurun_out/ct
timeout 600 python -m pytest tests/test_gpu_lr_residual.py tests/test_gpu_multiproc.py -x -q -k "consumers or residual" > gpurun_out/ct/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ct/pytest.log
tail -20 gpurun_out/ct/pytest.log | cut -c1-300
timeout 600 python bench.py --steps 20 --no-e2e 2>/dev/null | tail -1 | python -c '
import sys, json
d = json.loads(sys.stdin.read()); print(json.dumps(d["consumers"]["dense_lr"]))'
