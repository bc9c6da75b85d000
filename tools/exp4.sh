timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:wave_ -s 24 -c 8 -o gpurun_out/r1c_wave $B > gpurun_out/ncu3.log 2>&1
tail -3 gpurun_out/ncu3.log
