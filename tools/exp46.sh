# GPU parity tier (incl. device flatten, blit v2), output-stage timing v2 + ncu capture, ingestion timing
python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r1_pytest_gpu.log; cat gpurun_out/r1_pytest_gpu.log
timeout 120 python tools/bench_output.py > gpurun_out/r1_output_stage.json 2> gpurun_out/r1_output_stage.err; cut -c1-200 gpurun_out/r1_output_stage.json; tail -2 gpurun_out/r1_output_stage.err
timeout 120 python tools/bench_ingest.py > gpurun_out/r1_ingest.json 2> gpurun_out/r1_ingest.err; cat gpurun_out/r1_ingest.json; tail -2 gpurun_out/r1_ingest.err
timeout 200 ncu --set full --clock-control none --import-source on -k regex:blit_kernel -s 4 -c 2 -f -o gpurun_out/r1_blit python tools/bench_output.py > gpurun_out/r1_blit_ncu.log 2>&1; tail -1 gpurun_out/r1_blit_ncu.log
