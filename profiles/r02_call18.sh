# GPU call 18 (1 GPU): slot count of the node kernels' batch picked by the tile's largest degree (6 / 7 / 8)
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tracer.py tests/test_zz_no_alloc.py -x -q > gpurun_out/c18_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c18_pytest.log
timeout 600 python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c18_1m.json 2> gpurun_out/c18_1m.err; echo "1m exit $?"
timeout 900 python bench.py --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c18_5m.json 2> gpurun_out/c18_5m.err; echo "5m exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c18_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", round(d["ssh"]["ms_per_step"], 2), "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 3), d["roofline"]["traffic"])
        print("   ", {n: (round(v["ms"] / v["launches"], 3), v["launches"]) for n, v in k.items() if not n.startswith("k_amg") and not n.startswith("k_spmv") and v["ms"] / v["launches"] > 0.05})
    except Exception as e:
        print(f, "unreadable:", e)
PY
