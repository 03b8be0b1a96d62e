# GPU call 17 (1 GPU): launch-bound variant of k_update2 / k_fct_bounds2, then the ncu evidence of the final kernels at 1 M nodes
set -x
mkdir -p gpurun_out
B1="python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5"
timeout 600 $B1 > gpurun_out/c17_1m_def.json 2> gpurun_out/c17_1m_def.err; echo "def exit $?"
FB200_SO=fesom2_b200/csrc/libfesom_b200_varC.so timeout 600 $B1 > gpurun_out/c17_1m_varC.json 2> gpurun_out/c17_1m_varC.err; echo "varC exit $?"
FB200_SO=fesom2_b200/csrc/libfesom_b200_varC.so timeout 900 python bench.py --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c17_5m_varC.json 2> gpurun_out/c17_5m_varC.err; echo "5m varC exit $?"
N1="python bench.py --nodes 1000000 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
export FB200_SSH_GRAPH=0
timeout 600 $N1 > gpurun_out/c17_plain1.log 2>&1; echo "plain exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/c17_launches_1m.csv $N1 > gpurun_out/c17_ncu_list.log 2>&1; echo "list exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:^k_(ab2_g|edge_adv2|edge_diff2|ho_edge_fb|update2|lo2|fct_bounds2|impl_diff|grad_elem|clamp_salt)' -c 16 -o gpurun_out/prof_r02_tracer $N1 > gpurun_out/c17_ncu1.log 2>&1; echo "ncu tracer exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:^k_(spmv_alpha|spmv_sigma|x_update|amg_enter|amg_fold)' -s 30 -c 15 -o gpurun_out/prof_r02_ssh $N1 > gpurun_out/c17_ncu2.log 2>&1; echo "ncu ssh exit $?"
tail -n 2 gpurun_out/c17_ncu1.log gpurun_out/c17_ncu2.log
ls -la gpurun_out/*.ncu-rep
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c17_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", round(d["ssh"]["ms_per_step"], 2))
        print("   ", {n: (round(v["ms"] / v["launches"], 3), v["launches"]) for n, v in k.items() if not n.startswith("k_amg") and not n.startswith("k_spmv") and v["ms"] / v["launches"] > 0.05})
    except Exception as e:
        print(f, "unreadable:", e)
PY
