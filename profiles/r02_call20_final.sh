# GPU call 20 (1 GPU): the round's final state — whole GPU test suite, smoke(), the default bench line, the other BASELINE configs
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c20_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c20_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/c20_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/c20_smoke.log
timeout 1500 python bench.py > gpurun_out/c20_bench_default.json 2> gpurun_out/c20_bench_default.err; echo "default bench exit $?"
timeout 600 python bench.py --config core2like --no-e2e --no-cpu-baseline > gpurun_out/c20_bench_core2like.json 2> gpurun_out/c20_bench_core2like.err; echo "core2like exit $?"
timeout 600 python bench.py --config farclike --no-e2e --no-cpu-baseline > gpurun_out/c20_bench_farclike.json 2> gpurun_out/c20_bench_farclike.err; echo "farclike exit $?"
timeout 600 python bench.py --nodes 1000000 --shuffle --no-e2e --no-cpu-baseline > gpurun_out/c20_bench_1m_shuffled.json 2> gpurun_out/c20_bench_1m_shuffled.err; echo "shuffled exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c20_bench_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "value", round(d["value"] / 1e9, 3), "G  tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", {a: round(d["ssh"][a], 3) for a in ("ms_per_step", "iterations", "ms_per_iteration")},
              "e2e", (d.get("e2e") or {}).get("ms_per_step"), "cpu", (d.get("cpu_baseline") or {}).get("value"), "roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"], 3), "step", round(d["step_roofline"]["frac"], 3), "launches", d["gpu_launches"], "clocks", d["clocks"])
    except Exception as e:
        print(f, "unreadable:", e)
PY
for f in gpurun_out/c20_*.err; do echo $f; tail -n 2 $f; done
