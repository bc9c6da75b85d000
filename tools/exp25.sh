B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
for c in 512 256 128 64 32 16; do echo "== cap $c"; VRT_ROUND_CAP=$c $B | python -c "$J"; done
