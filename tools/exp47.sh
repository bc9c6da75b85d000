# blit v3: parity tests of the output stage, timing, ncu capture
python -m pytest tests -m gpu -q -k "blit or window or flatten or cpp" 2>&1 | tail -4
timeout 120 python tools/bench_output.py > gpurun_out/r1_output_stage.json 2> gpurun_out/r1_output_stage.err; cut -c1-200 gpurun_out/r1_output_stage.json; tail -2 gpurun_out/r1_output_stage.err
timeout 200 ncu --set full --clock-control none --import-source on -k regex:blit_kernel -s 4 -c 2 -f -o gpurun_out/r1_blit python tools/bench_output.py > gpurun_out/r1_blit_ncu.log 2>&1; tail -1 gpurun_out/r1_blit_ncu.log
