# GPU call 4: zero-entry sums + 16-byte record loads; launch-bound variant U; FMA-contracted build (speed and deviation)
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_tracer.py -x -q > gpurun_out/c4_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c4_pytest.log
B="python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5"
C=$PWD/fesom2_b200/csrc
$B > gpurun_out/c4_def.json 2> gpurun_out/c4_def.err; echo "def exit $?"
FB200_SO=$C/libvarU.so $B > gpurun_out/c4_varU.json 2> gpurun_out/c4_varU.err; echo "U exit $?"
FB200_SO=$C/libfmad.so $B > gpurun_out/c4_fmad.json 2> gpurun_out/c4_fmad.err; echo "fmad exit $?"
FB200_SO=$C/libfmad.so python -m pytest tests/test_gpu_tracer.py -q -k "test_tracer_step_matches_oracle" > gpurun_out/c4_pytest_fmad.log 2>&1; echo "pytest fmad exit $?"; tail -5 gpurun_out/c4_pytest_fmad.log
python bench.py --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c4_def_5m.json 2> gpurun_out/c4_def_5m.err; echo "5m exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c4_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2),
              {n: (round(v["ms"] / v["launches"], 3), v["launches"]) for n, v in k.items() if not n.startswith("k_amg") and not n.startswith("k_spmv") and n not in ("k_x_update","k_init_res","k_scale_rows")})
    except Exception as e:
        print(f, "unreadable:", e)
PY
