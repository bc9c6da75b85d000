B="python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d["roofline"]; print(round(d["ms_per_step"],3), r.get("stages_ms")["anyhit"], r.get("stages_Mrays")["anyhit"])'
echo "== AO only, engine"; $B --passes 29 | python -c "$J"
echo "== AO only, old"; VRT_ENGINE_T=0 $B --passes 29 | python -c "$J"
echo "== shadow only, engine"; $B --passes 27 | python -c "$J"
echo "== shadow only, old"; VRT_ENGINE_T=0 $B --passes 27 | python -c "$J"
echo "== both, engine"; $B --passes 31 | python -c "$J"
echo "== both, old"; VRT_ENGINE_T=0 $B --passes 31 | python -c "$J"
