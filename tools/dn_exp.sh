# usage: tools/dn_exp.sh "<suffixes>": per-launch times of the denoiser kernels of a cfg3 frame for build variants (ncu launch list)
for v in $1; do [ "$v" = "-" ] && v=""; echo "== libvrt$v"
VRT_LIB=$PWD/voxel-raytracing_b200/libvrt$v.so ncu --metrics gpu__time_duration.sum --clock-control none -k regex:denoise -s 6 -c 4 python bench.py --config cfg3 --steps 2 --warmup 3 --min-timed-ms 20 --no-cpu-baseline 2>&1 | grep -E "gpu__time" | tr -s ' ' | paste -sd' '
done
