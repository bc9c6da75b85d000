# OCTANT traversal (directional safe-step fields): parity, then bench per traversal
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
for t in 4 5; do echo "== traversal $t"; $B --traversal $t | python -c "$J"; done
echo "== traversal 5 ao only"; $B --traversal 5 --passes 21 | python -c "$J"
echo "== traversal 5 shadow only"; $B --traversal 5 --passes 19 | python -c "$J"
echo "== traversal 5 engine mode2 T8"; VRT_ENGINE_T=8 $B --traversal 4 | python -c "$J"
