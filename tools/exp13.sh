B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --traversal 5"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
for v in 0 1 2 3 4; do for f in 8 6 4 2; do echo "== variant $v farshare $f"; VRT_ENGINE_FARSHARE=$f VRT_ENGINE_VARIANT=$v VRT_ENGINE_T=8 $B | python -c "$J"; done; done
