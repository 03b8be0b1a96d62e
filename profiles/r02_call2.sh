set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_tracer.py -x -q -k "matches_oracle" > gpurun_out/c2_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c2_pytest.log
B="python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5"
$B > gpurun_out/c2_v2.json 2> gpurun_out/c2_v2.err; echo "v2 exit $?"
FB200_NT1=1 $B > gpurun_out/c2_v2nt1.json 2> gpurun_out/c2_v2nt1.err; echo "nt1 exit $?"
FB200_V1=all $B > gpurun_out/c2_v1.json 2> gpurun_out/c2_v1.err; echo "v1 exit $?"
python bench.py --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c2_v2_5m.json 2> gpurun_out/c2_v2_5m.err; echo "5m exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c2_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh_ms", round(d["ssh"]["ms_per_step"], 2),
              {n: (round(v["ms"] / v["launches"], 3), v["launches"]) for n, v in k.items() if not n.startswith("k_amg") and not n.startswith("k_spmv") and n not in ("k_x_update","k_init_res","k_scale_rows")})
    except Exception as e:
        print(f, "unreadable:", e)
PY
tail -3 gpurun_out/c2_*.err
