# where does the any-hit wave's time go? old run-to-completion kernel, AO-only vs shadow-only vs both (reflections off / on)
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), r.get("stages_ms"), r.get("stages_Mrays"), r.get("stages_Gvisits"))'
for p in 19 21 23 27 29 31; do echo "== passes $p (1 primary 2 shadow 4 ao 8 reflect 16 shade)"; VRT_ENGINE_T=0 $B --passes $p | python -c "$J"; done
echo "== ao 1"; VRT_ENGINE_T=0 $B --passes 21 --ao 1 | python -c "$J"
