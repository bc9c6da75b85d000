# usage: tools/profile_round.sh <round-tag>   (on the GPU box, via gpurun). Produces, under gpurun_out/:
#   <tag>_bench.json          the plain bench line (never under a profiler)
#   <tag>_launches.csv        ncu launch list of the same command (gpu__time_duration.sum per launch; cold-cache, serialised)
#   <tag>_full.ncu-rep        one --set full capture of the frame's kernels (primary, 5 reflection waves, any-hit engine, resolve)
T=$1
B="python bench.py --steps 3 --warmup 3"
$B > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || exit 1
K='regex:wave_|render_kernel|denoise_kernel|detile_kernel'
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 60 --csv --log-file gpurun_out/${T}_launches.csv $B --no-cpu-baseline > gpurun_out/${T}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k "$K" -s 32 -c 8 -f -o gpurun_out/${T}_full $B --no-cpu-baseline > gpurun_out/${T}_ncu2.log 2>&1
tail -2 gpurun_out/${T}_ncu2.log; cat gpurun_out/${T}_bench.json | cut -c1-300
