# closing verification of the tree as committed: GPU parity tier, smoke, default bench line
python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r1_pytest_gpu.log; cat gpurun_out/r1_pytest_gpu.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1_smoke.log 2>&1; tail -1 gpurun_out/r1_smoke.log
timeout 200 python bench.py > gpurun_out/r1_bench_cfg4.json 2> gpurun_out/r1_bench.err; cut -c1-260 gpurun_out/r1_bench_cfg4.json; tail -2 gpurun_out/r1_bench.err
