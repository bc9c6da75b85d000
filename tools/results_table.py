"""Markdown rows of BASELINE.md section 4 from the bench lines committed under profiles/ (usage: results_table.py)."""
import json, os
R = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles")


def line(name):
    p = os.path.join(R, name)
    return json.loads(open(p).read().strip().splitlines()[-1]) if os.path.exists(p) else None


import sys
T = sys.argv[1] if len(sys.argv) > 1 else "r2"      # round tag of the files under profiles/
rows = [("1", "treehouse stand-in", f"{T}_bench_cfg1.json", "1"), ("2", "mandelbulb stand-in", f"{T}_bench_cfg2.json", "1"),
        ("3", "floatingcolored stand-in", f"{T}_bench_cfg3.json", "1"), ("3", "floatingcolored stand-in (denoiser distributed: strips)", f"scale/{T}_scale_cfg3_N2.json", "2"),
        ("3", "floatingcolored stand-in (denoiser distributed: strips)", f"scale/{T}_scale_cfg3_N4.json", "4"),
        ("3", "floatingcolored stand-in (denoiser distributed: strips)", f"scale/{T}_scale_cfg3_N8.json", "8"), ("4", "synthetic 1024^3 terrain", f"{T}_bench_cfg4.json", "1"),
        ("4", "synthetic 1024^3 terrain", f"scale/{T}_scale_cfg4_N2.json", "2"), ("4", "synthetic 1024^3 terrain", f"scale/{T}_scale_cfg4_N4.json", "4"),
        ("4", "synthetic 1024^3 terrain", f"scale/{T}_scale_cfg4_N8.json", "8"), ("5", "synthetic 2048^3 sparse", f"{T}_bench_cfg5_1gpu.json", "1"),
        ("5", "synthetic 2048^3 sparse, orbit", f"scale/{T}_scale_cfg5_N8.json", "8")]
print("| Config | Scene | Resolution | GPUs | ms/frame | e2e ms/frame | Mrays/s (all) | Mrays/s (primary) | steps_ref / frame | bytes_alg / frame | roofline (dominant kernel vs HBM; issue; primary vs L2) | CPU oracle Mrays/s (cores) | parity vs oracle |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
for cfg, scene, f, n in rows:
    d = line(f)
    if not d:
        continue
    r = d["roofline"]; cb = d.get("cpu_baseline")
    print(f"| {cfg} | {scene} | {d['config']['resolution'][0]}x{d['config']['resolution'][1]} | {n} | {d['ms_per_step']:.3f} | {d['e2e']['ms_per_step']:.3f} | "
          f"{d['value']:.0f} | {d['Mrays_per_s_primary']:.0f} | {d['voxel_visits_per_frame']:.3g} | {r['frame']['algorithmic_bytes']:.3g} | "
          f"{r['kernel'].split(' ')[0]} {r['frac']:.2f}; issue {(r.get('issue') or {}).get('frac', 0):.2f}; L2 {(r.get('l2') or {}).get('frac', 0):.2f} | "
          + (f"{cb['value']:.1f} ({cb['cores']})" if cb else "-") + " | "
          + (("0 mismatches on %d px" % d["parity"]["pixels_compared"]) if d.get("parity", {}).get("ok") else ("-" if "parity" not in d else "MISMATCH")) + " |")
