# LINEAR vs FINE traversal: parity tests, then bench per traversal (wavefront), stage split
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r["kernel"], r.get("stages_ms"), d["gpu_launches"])'
for t in 3 4; do echo "== traversal $t"; $B --traversal $t | python -c "$J"; done
for t in 3 4; do echo "== traversal $t primary only"; $B --traversal $t --passes 1 | python -c "$J"; done
