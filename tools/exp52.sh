python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/r1_pytest_gpu.log; cat gpurun_out/r1_pytest_gpu.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1_smoke.log 2>&1; tail -1 gpurun_out/r1_smoke.log
