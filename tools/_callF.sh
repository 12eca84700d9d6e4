mkdir -p gpurun_out/cf
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/cf/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/cf/pytest.log
tail -6 gpurun_out/cf/pytest.log | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/cf/bench1.log 2>gpurun_out/cf/bench1.err; echo "bench rc=$?"; tail -1 gpurun_out/cf/bench1.log | cut -c1-400
