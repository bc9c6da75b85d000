timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
echo "== cfg3"; $B --config cfg3 | python -c "$J"
bash tools/launches.sh cfg3_launches --config cfg3 2>&1 | grep denoise | head -2
