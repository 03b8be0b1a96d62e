# First GPU call of the next round (run as: gpurun --timeout 2400 -- 'bash profiles/r02_call_first.sh'): everything that was
# written after round 1's GPU budget was spent gets its first run on a B200, and the prepared experiments are measured.
#  1. the whole GPU suite, including tests/test_gpu_zz_*.py (relax_zonal_temp, output means, SSH assembly; multi-rank ALE is
#     skipped on one GPU) — all of them pass under the CPU emulation (tests/test_emulated*.py);
#  2. the default bench line at 5 M nodes with --e2e-ale (device-side ALE producers end to end; 1 M was 110.6 -> 77.6 ms);
#  3. the same kernels with FB200_EDGE_ORDER=tile at 1 M and 5 M nodes (A/B of the internal edge numbering, no e2e / CPU legs);
#  4. the launch list of the default command for profiles/.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest.log; tail -5 gpurun_out/r02_pytest.log
python bench.py --e2e-ale > gpurun_out/r02_bench_5m_ale.json 2> gpurun_out/r02_bench_5m_ale.err; echo "bench exit $?"
for n in 1000000 5000000; do
  python bench.py --nodes $n --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/r02_ab_ref_$n.json 2> gpurun_out/r02_ab_ref_$n.err; echo "A exit $?"
  FB200_EDGE_ORDER=tile python bench.py --nodes $n --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/r02_ab_tile_$n.json 2> gpurun_out/r02_ab_tile_$n.err; echo "B exit $?"
done
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
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches.csv python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/r02_ncu.log 2>&1; echo "ncu exit $?"
