# denoiser: edge-blind kernel for iteration 0 + host-side tap offsets; parity tier and the cfg3 line
python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/r1_pytest_gpu.log; cat gpurun_out/r1_pytest_gpu.log
timeout 120 python bench.py --config cfg3 --steps 20 --warmup 3 > gpurun_out/r1_bench_cfg3.json 2> gpurun_out/r1_bench3.err; tail -2 gpurun_out/r1_bench3.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r1_bench_cfg3.json").read().strip().splitlines()[-1]); print("cfg3 ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), {k: (round(v,3) if isinstance(v,float) else v) for k,v in d["roofline"].get("stages_ms",{}).items()}, [k for k in d.keys()][:40])
PY
