# usage: tools/launches.sh <out-name> [bench args]: per-launch durations of the frame kernels (ncu serialises: shares, not absolutes)
O=$1; shift
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"wave_|render_kernel|denoise|detile" -s 33 -c 44 --csv --log-file gpurun_out/${O}.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline $@ > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${O}.csv")) if len(r)>10]
h=rows[0]; ik=h.index("Kernel Name"); iv=h.index("Metric Value")
for r in rows[1:]: print(r[ik][:70].ljust(70), r[iv])
PY
