timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
echo "== default"; $B | python -c "$J"
for c in cfg1 cfg2 cfg3; do echo "== $c"; $B --config $c | python -c "$J"; done
python - <<'PY'
import sys, time
sys.path.insert(0, "voxel-raytracing_b200")
import pyvrt
ctx = pyvrt.Context(0)
t0 = time.perf_counter(); gs = pyvrt.VoxelScene.generate(ctx, 1, 42, 1024, 1024, 1024); ctx.sync(); t1 = time.perf_counter()
gs.download_odist(0)[0, 0, 0]; t2 = time.perf_counter()
print("scene generate + brick/fine levels: %.1f ms; octant fields (8 x 1 GiB): %.1f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3))
PY
