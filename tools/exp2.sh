# FINE traversal: parity tests, then bench vs BRICK, pass splits
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
J='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), round(d["value"],1), d["rays_per_frame"], d["voxel_visits_per_frame"])'
for t in 2 3; do echo "== traversal $t all passes"; $B --traversal $t | python -c "$J"; done
for pm in 1 19 21 23; do echo "== traversal 3 passes $pm"; $B --traversal 3 --passes $pm | python -c "$J"; done
for c in cfg1 cfg2 cfg3; do echo "== $c"; $B --config $c | python -c "$J"; done
