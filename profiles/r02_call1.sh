# GPU call 1 of round 2: new tests (Redi explicit term, biharmonic filter, allocation-free extrema) + everything that had only run
# under the emulation, the edge-order A/B at 1 M nodes, the 5 M default line with --e2e-ale, and a launch list.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv
python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r02_noalloc_alone.log 2>&1; echo "noalloc alone exit $?" | tee -a gpurun_out/r02_noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r02_pytest.log; tail -5 gpurun_out/r02_pytest.log
for n in 1000000; do
  python bench.py --nodes $n --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/r02_ab_ref_$n.json 2> gpurun_out/r02_ab_ref_$n.err; echo "A exit $?"
  FB200_EDGE_ORDER=tile python bench.py --nodes $n --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/r02_ab_tile_$n.json 2> gpurun_out/r02_ab_tile_$n.err; echo "B exit $?"
done
python bench.py --e2e-ale --no-cpu-baseline --steps 5 > gpurun_out/r02_bench_5m_ale.json 2> gpurun_out/r02_bench_5m_ale.err; echo "bench exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh_ms", round(d["ssh"]["ms_per_step"], 2), "e2e", d.get("e2e", {}).get("ms_per_step"),
              "e2e_ale", d.get("e2e_device_producers", {}).get("ms_per_step"),
              {n: round(v["ms"] / v["launches"], 3) for n, v in k.items() if n in ("k_update", "k_ho_edge", "k_fct_bounds", "k_lo", "k_diff_edge")})
    except Exception as e:
        print(f, "unreadable:", e)
PY
