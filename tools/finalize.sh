# usage: tools/finalize.sh <tag>  — the round's single-GPU measurement batch (on the GPU box via gpurun): GPU test tier, smoke, the bench
# line of every config, the reference arm, the traversal back-ends and pipelines kept for measurement. Output: gpurun_out/<tag>_*.
T=$1
(time python -m pytest tests -m gpu -q --durations=8) > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -14 gpurun_out/${T}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_ref.err
python bench.py > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench.err
for c in cfg1 cfg2 cfg3; do python bench.py --config $c --steps 20 --warmup 3 > gpurun_out/${T}_bench_$c.json 2>> gpurun_out/${T}_bench.err; done
python bench.py --config cfg5 --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_cfg5_1gpu.json 2>> gpurun_out/${T}_bench.err
for t in 1 2 3 4 6; do python bench.py --traversal $t --steps 5 --warmup 3 --no-cpu-baseline --min-timed-ms 300 > gpurun_out/${T}_bench_cfg4_traversal$t.json 2>> gpurun_out/${T}_bench.err; done
python bench.py --pipeline 1 --steps 5 --warmup 3 --no-cpu-baseline --min-timed-ms 300 > gpurun_out/${T}_bench_cfg4_megakernel.json 2>> gpurun_out/${T}_bench.err
python bench.py --passes 1 --steps 10 --warmup 3 --no-cpu-baseline --min-timed-ms 300 > gpurun_out/${T}_bench_cfg4_primary_only.json 2>> gpurun_out/${T}_bench.err
for f in gpurun_out/${T}_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(sys.argv[1].split("/")[-1].ljust(44), "ms", round(d["ms_per_step"], 3), "Mrays/s", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "cpu", (d.get("cpu_baseline") or {}).get("value"), "frac", r.get("frac"), "l2", (r.get("l2") or {}).get("frac"), "parity", (d.get("parity") or {}).get("ok"))
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
tail -3 gpurun_out/${T}_bench.err
