# GPU call 15 (2 GPUs): overlap exchange in the head of the cycle's first kernel, last level collapsed into one matrix
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_ssh.py -x -q > gpurun_out/c15_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c15_pytest.log
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
B="bench.py --gpus 2 --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5"
$T $B > gpurun_out/c15_1m.json 2> gpurun_out/c15_1m.err; echo "1m exit $?"
$T bench.py --gpus 2 --no-cpu-baseline --no-e2e --no-parity --steps 5 > gpurun_out/c15_5m.json 2> gpurun_out/c15_5m.err; echo "5m exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c15_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", {a: d["ssh"][a] for a in ("ms_per_step", "iterations", "ms_per_iteration")}, "parity", (d.get("parity") or {}).get("ok"))
        print("   ", {n: (round(v["ms"] / v["launches"], 4), v["launches"], v.get("GBps")) for n, v in k.items() if n.startswith("k_amg") or n.startswith("k_spmv") or n.startswith("k_overlap") or n in ("k_x_update",)})
    except Exception as e:
        print(f, "unreadable:", e)
PY
for f in gpurun_out/c15_*.err; do echo $f; tail -n 3 $f; done
