cd voxel-raytracing_b200
B="python ../bench.py --steps 5 --warmup 3 --no-cpu-baseline"
for b in 5 6 8; do echo "== minblocks $b"; VRT_LIB=$PWD/libvrt_b$b.so $B | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'])"; done
echo "== pass splits (default lib)"
for pm in 1 17 19 21 23 31; do echo "passes $pm"; $B --passes $pm | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['rays_per_frame'], d['voxel_visits_per_frame'])"; done
echo "== no gbuffer"; $B --no-gbuffer | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'])"
echo "== cfg1 cfg2 cfg3"
for c in cfg1 cfg2 cfg3; do $B --config $c | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['config']['workload'][:30], d['ms_per_step'], d['value'], d['rays_per_frame'], d['voxel_visits_per_frame'])"; done
