# refill engine for the any-hit wave: parity, then threshold sweep
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --traversal 4"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
for t in 0 1 4 8 12 16 24; do echo "== VRT_ENGINE_T $t"; VRT_ENGINE_T=$t $B | python -c "$J"; done
