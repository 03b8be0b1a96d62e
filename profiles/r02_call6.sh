set -x
mkdir -p gpurun_out
python bench.py --steps 5 --cpu-steps 3 > gpurun_out/c6_full_5m.json 2> gpurun_out/c6_full_5m.err; echo "5m exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c6_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", round(d["ssh"]["ms_per_step"],2), "e2e", d.get("e2e",{}).get("ms_per_step"), "cpu", {a: d.get("cpu_baseline",{}).get(a) for a in ("value","cores","tracer_ms_per_step","ssh_ms_per_step","steps_timed","sample_surface_nodes_total")},
              {n: (round(v["ms"] / v["launches"], 3), v["launches"], v["GBps"]) for n, v in k.items() if not n.startswith("k_amg") and not n.startswith("k_spmv") and n not in ("k_x_update","k_init_res","k_scale_rows")})
        print(" roofline", d.get("roofline"), d.get("step_roofline"))
    except Exception as e:
        print(f, "unreadable:", e)
PY
tail -n 5 gpurun_out/c6_full_5m.err
