# GPU call 3: column kernels with batched loads — launch-bound variants and tracers-per-launch A/B at 1 M nodes, then 5 M
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_tracer.py -x -q -k "matches_oracle" > gpurun_out/c3_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c3_pytest.log
B="python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5"
C=$PWD/fesom2_b200/csrc
$B > gpurun_out/c3_def.json 2> gpurun_out/c3_def.err; echo "def exit $?"
FB200_SO=$C/libvarX.so $B > gpurun_out/c3_varX.json 2> gpurun_out/c3_varX.err; echo "X exit $?"
FB200_SO=$C/libvarY.so $B > gpurun_out/c3_varY.json 2> gpurun_out/c3_varY.err; echo "Y exit $?"
FB200_SO=$C/libvarZ.so $B > gpurun_out/c3_varZ.json 2> gpurun_out/c3_varZ.err; echo "Z exit $?"
FB200_NT_LO=1 FB200_NT_EDGE=1 FB200_NT_DIFF=1 FB200_NT_GRAD=1 $B > gpurun_out/c3_nt1.json 2> gpurun_out/c3_nt1.err; echo "nt1 exit $?"
FB200_NT_BOUNDS=2 FB200_NT_UPDATE=2 $B > gpurun_out/c3_nt2.json 2> gpurun_out/c3_nt2.err; echo "nt2 exit $?"
python bench.py --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c3_def_5m.json 2> gpurun_out/c3_def_5m.err; echo "5m exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c3_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2),
              {n: (round(v["ms"] / v["launches"], 3), v["launches"]) for n, v in k.items() if not n.startswith("k_amg") and not n.startswith("k_spmv") and n not in ("k_x_update","k_init_res","k_scale_rows")})
    except Exception as e:
        print(f, "unreadable:", e)
PY
for f in gpurun_out/c3_*.err; do echo $f; tail -n 2 $f; done
