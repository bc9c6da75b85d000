# engine v3 (fixed-length predicated rounds, recompute look-ups): parity with the engine ON, then threshold sweep
VRT_ENGINE_T=8 timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
for tr in 5 4; do for t in 0 4 8 16; do echo "== traversal $tr T $t"; VRT_ENGINE_T=$t $B --traversal $tr | python -c "$J"; done; done
echo "== traversal 5 ao only T8"; VRT_ENGINE_T=8 $B --traversal 5 --passes 21 | python -c "$J"
echo "== traversal 5 shadow only T8"; VRT_ENGINE_T=8 $B --traversal 5 --passes 19 | python -c "$J"
