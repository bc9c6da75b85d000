# wavefront pipeline: parity tests, then bench megakernel vs wavefront, stage split
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r["kernel"], r.get("stages_ms"), r.get("stages_Mrays"), r.get("stages_Gvisits"), d["gpu_launches"])'
for p in 1 2; do echo "== pipeline $p"; $B --pipeline $p | python -c "$J"; done
for c in cfg1 cfg2 cfg3; do echo "== $c"; $B --config $c | python -c "$J"; done
