# flat lock-step engine for the any-hit wave: parity, then round-mode x threshold sweep
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), round(d["value"],1), r.get("stages_ms"))'
echo "== VRT_ENGINE_T 0"; VRT_ENGINE_T=0 $B | python -c "$J"
for m in 0 1 2; do for t in 4 8 16; do echo "== MODE $m T $t"; VRT_ENGINE_MODE=$m VRT_ENGINE_T=$t $B | python -c "$J"; done; done
