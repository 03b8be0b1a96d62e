# GPU call 5: shared / constant reciprocals (exact-division sequence) in k_lo2, k_edge_adv2, k_impl_diff; new bench line incl. the
# device-producer e2e leg and the cpu_baseline leg with the timing builds
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_tracer.py -x -q > gpurun_out/c5_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c5_pytest.log
python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c5_def.json 2> gpurun_out/c5_def.err; echo "def exit $?"
python bench.py --steps 5 --cpu-steps 3 > gpurun_out/c5_full_5m.json 2> gpurun_out/c5_full_5m.err; echo "5m exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c5_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", round(d["ssh"]["ms_per_step"],2), "e2e", d.get("e2e",{}).get("ms_per_step"), "cpu", {a: d.get("cpu_baseline",{}).get(a) for a in ("value","cores","tracer_ms_per_step","ssh_ms_per_step","steps_timed")},
              {n: (round(v["ms"] / v["launches"], 3), v["launches"], v["GBps"]) for n, v in k.items() if not n.startswith("k_amg") and not n.startswith("k_spmv") and n not in ("k_x_update","k_init_res","k_scale_rows")})
        print(" roofline", d.get("roofline"), d.get("step_roofline"))
    except Exception as e:
        print(f, "unreadable:", e)
PY
tail -n 5 gpurun_out/c5_full_5m.err
