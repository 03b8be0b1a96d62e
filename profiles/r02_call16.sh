# GPU call 16 (1 GPU): packed edge / node fields (one 16-byte gather per slot in k_lo2, k_fct_bounds2, k_update2)
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tracer.py tests/test_zz_no_alloc.py -x -q > gpurun_out/c16_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c16_pytest.log
B="timeout 600 python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5"
run() { name=$1; shift; env "$@" $B > gpurun_out/c16_$name.json 2> gpurun_out/c16_$name.err; echo "$name exit $?"; }
run def A=1
run lo1 FB200_NT_LO=1
run b2 FB200_NT_BOUNDS=2
run varA FB200_SO=fesom2_b200/csrc/libfesom_b200_varA.so
run varB FB200_SO=fesom2_b200/csrc/libfesom_b200_varB.so
run varB_lo1 FB200_SO=fesom2_b200/csrc/libfesom_b200_varB.so FB200_NT_LO=1
python - <<'PY' > gpurun_out/c16_best.env
import json, glob
best = None
envs = {"def": "", "lo1": "FB200_NT_LO=1", "b2": "FB200_NT_BOUNDS=2", "varA": "FB200_SO=fesom2_b200/csrc/libfesom_b200_varA.so",
        "varB": "FB200_SO=fesom2_b200/csrc/libfesom_b200_varB.so", "varB_lo1": "FB200_SO=fesom2_b200/csrc/libfesom_b200_varB.so FB200_NT_LO=1"}
for f in sorted(glob.glob("gpurun_out/c16_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception:
        continue
    nm = f.split("c16_")[1][:-5]
    if nm in envs and (best is None or d["tracer_ms_per_step"] < best[0]):
        best = (d["tracer_ms_per_step"], nm)
print(envs[best[1]] if best else "")
PY
cat gpurun_out/c16_best.env
env $(cat gpurun_out/c16_best.env) timeout 900 python bench.py --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c16_5m.json 2> gpurun_out/c16_5m.err; echo "5m exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c16_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", round(d["ssh"]["ms_per_step"], 2))
        print("   ", {n: (round(v["ms"] / v["launches"], 3), v["launches"]) for n, v in k.items() if not n.startswith("k_amg") and not n.startswith("k_spmv") and v["ms"] / v["launches"] > 0.05})
    except Exception as e:
        print(f, "unreadable:", e)
PY
for f in gpurun_out/c16_*.err; do echo $f; tail -n 2 $f; done
