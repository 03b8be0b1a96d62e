# GPU call 13 (2 GPUs): interface pushes fused into the consuming kernels, 4 threads per row on the coarse levels, SSH system resident
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_ssh.py -x -q > gpurun_out/c13_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c13_pytest.log
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
B="bench.py --gpus 2 --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5"
$T $B > gpurun_out/c13_1m.json 2> gpurun_out/c13_1m.err; echo "1m exit $?"
FB200_SSH_COMM=nccl $T $B --no-parity > gpurun_out/c13_1m_nccl.json 2> gpurun_out/c13_1m_nccl.err; echo "nccl exit $?"
$T bench.py --gpus 2 --no-cpu-baseline --steps 5 > gpurun_out/c13_5m.json 2> gpurun_out/c13_5m.err; echo "5m exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c13_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", {a: d["ssh"][a] for a in ("ms_per_step", "iterations", "ms_per_iteration")}, "e2e", d.get("e2e", {}).get("ms_per_step"), "parity", (d.get("parity") or {}).get("ok"))
        print("   ", {n: (round(v["ms"] / v["launches"], 4), v["launches"]) for n, v in k.items() if n.startswith("k_amg") or n.startswith("k_spmv") or n.startswith("nccl") or n.startswith("k_overlap") or n in ("k_x_update","k_apply_scalars","halo_exchange_flat","overlap_exchange")})
    except Exception as e:
        print(f, "unreadable:", e)
PY
for f in gpurun_out/c13_*.err; do echo $f; tail -n 3 $f; done
