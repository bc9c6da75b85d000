# usage: tools/profile_round.sh <round-tag>   (on the GPU box, via gpurun). Produces, under gpurun_out/:
#   <tag>_bench.json          the plain bench line (never under a profiler)
#   <tag>_launches.csv        ncu launch list of the same command (gpu__time_duration.sum per launch; cold-cache, serialised)
#   <tag>_full.ncu-rep        one --set full capture of the frame's kernels (primary, reflection waves, any-hit engine, resolve)
#   <tag>_denoise.ncu-rep     the same for the two denoiser launches of a cfg3 frame
T=$1
B="python bench.py --steps 3 --warmup 3 --min-timed-ms 50"
$B > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || exit 1
K='regex:wave_|render_kernel|denoise_kernel|push_tiles'
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 60 --csv --log-file gpurun_out/${T}_launches.csv $B --no-cpu-baseline > gpurun_out/${T}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k "$K" -s 36 -c 6 -f -o gpurun_out/${T}_full $B --no-cpu-baseline > gpurun_out/${T}_ncu2.log 2>&1
tail -2 gpurun_out/${T}_ncu2.log
python bench.py --config cfg3 --steps 3 --warmup 3 --min-timed-ms 50 --no-cpu-baseline > gpurun_out/${T}_profrun_cfg3.json 2>> gpurun_out/${T}_bench.err
ncu --set full --clock-control none --import-source on -k regex:denoise_kernel -s 8 -c 2 -f -o gpurun_out/${T}_denoise python bench.py --config cfg3 --steps 3 --warmup 3 --min-timed-ms 50 --no-cpu-baseline > gpurun_out/${T}_ncu3.log 2>&1
tail -2 gpurun_out/${T}_ncu3.log; cut -c1-300 gpurun_out/${T}_bench.json
