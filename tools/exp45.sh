# round-1 closing batch (single GPU, short): GPU parity tier incl. the output stage, output-stage timing, refreshed default bench line
python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/r1_pytest_gpu.log; cat gpurun_out/r1_pytest_gpu.log
timeout 120 python tools/bench_output.py > gpurun_out/r1_output_stage.json 2> gpurun_out/r1_output_stage.err; cat gpurun_out/r1_output_stage.json | cut -c1-260; tail -2 gpurun_out/r1_output_stage.err
timeout 200 python bench.py > gpurun_out/r1_bench_cfg4.json 2> gpurun_out/r1_bench.err; cut -c1-400 gpurun_out/r1_bench_cfg4.json; tail -2 gpurun_out/r1_bench.err
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1_smoke.log 2>&1; tail -1 gpurun_out/r1_smoke.log
timeout 200 ncu --set full --clock-control none --import-source on -k regex:blit_kernel -s 4 -c 2 -f -o gpurun_out/r1_blit python tools/bench_output.py > gpurun_out/r1_blit_ncu.log 2>&1; tail -2 gpurun_out/r1_blit_ncu.log
