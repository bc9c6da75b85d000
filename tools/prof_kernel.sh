# usage: tools/prof_kernel.sh <kernel-regex> <out-name> [bench args...]  — one ncu --set full capture of a kernel (after the plain run exits 0)
K=$1; O=$2; shift 2
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline $@"
$B > gpurun_out/${O}_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$K -s 6 -c 2 -o gpurun_out/$O -f $B > gpurun_out/${O}_ncu.log 2>&1
tail -2 gpurun_out/${O}_ncu.log
