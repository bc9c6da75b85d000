tools/l2peak 64 > gpurun_out/l2peak.json; cat gpurun_out/l2peak.json
echo "== cfg5 on one GPU (2048^3 sparse, 7680x4320)"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"), d["rays_per_frame"], d["voxel_visits_per_frame"], d["e2e"])'
(nvidia-smi --query-gpu=memory.used --format=csv,noheader -lms 500 > gpurun_out/cfg5_mem.log &) 
SECONDS=0
python bench.py --config cfg5 --steps 4 --warmup 3 --no-cpu-baseline 2> gpurun_out/cfg5.err | python -c "$J"
echo "wall $SECONDS s"; tail -3 gpurun_out/cfg5.err
sort -n gpurun_out/cfg5_mem.log | tail -1
