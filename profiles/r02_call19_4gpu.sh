# GPU call 19 (4 GPUs): strong scaling of the 5 M-node step on 4 METIS partitions, partitioned-vs-undivided parity on 4 ranks
set -x
mkdir -p gpurun_out
T="timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29611"
$T bench.py --gpus 4 --no-cpu-baseline --no-e2e --steps 5 > gpurun_out/c19_5m_4gpu.json 2> gpurun_out/c19_5m_4gpu.err; echo "5m 4gpu exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c19_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", {a: d["ssh"][a] for a in ("ms_per_step", "iterations", "ms_per_iteration")}, "parity", d.get("parity"))
        print("   ", {n: (round(v["ms"] / v["launches"], 4), v["launches"]) for n, v in k.items() if n.startswith("k_amg") or n.startswith("k_spmv") or n.startswith("k_overlap") or n.startswith("nccl") or n.startswith("k_halo") or n in ("k_x_update",)})
    except Exception as e:
        print(f, "unreadable:", e)
PY
tail -n 5 gpurun_out/c19_5m_4gpu.err
