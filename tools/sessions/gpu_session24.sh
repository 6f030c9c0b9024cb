mkdir -p gpurun_out/s24
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s24/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s24/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/s24/smoke.log 2>&1
