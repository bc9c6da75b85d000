timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
echo "== default"; $B | python -c "$J"
echo "== engine off"; VRT_ENGINE_T=0 $B | python -c "$J"
for c in cfg1 cfg2 cfg3; do echo "== $c"; $B --config $c | python -c "$J"; done
