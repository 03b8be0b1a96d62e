# GPU call 7 (2 GPUs): multi-GPU parity tests (first driver-visible logs), peer-memory transport vs NCCL in the SSH solve,
# fused coarse levels, bench parity block
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_zz_multi_ale.py -x -q > gpurun_out/c7_pytest_multi.log 2>&1; echo "pytest multi exit $?"; tail -4 gpurun_out/c7_pytest_multi.log
timeout 300 python -m pytest tests/test_gpu_ssh.py -x -q > gpurun_out/c7_pytest_ssh.log 2>&1; echo "pytest ssh exit $?"; tail -3 gpurun_out/c7_pytest_ssh.log
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
$T bench.py --gpus 2 --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c7_p2p.json 2> gpurun_out/c7_p2p.err; echo "p2p exit $?"
FB200_SSH_COMM=nccl $T bench.py --gpus 2 --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5 --no-parity > gpurun_out/c7_nccl.json 2> gpurun_out/c7_nccl.err; echo "nccl exit $?"
FB200_AMG_FUSED=0 $T bench.py --gpus 2 --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5 --no-parity > gpurun_out/c7_unfused.json 2> gpurun_out/c7_unfused.err; echo "unfused exit $?"
timeout 300 python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c7_1gpu.json 2> gpurun_out/c7_1gpu.err; echo "1gpu exit $?"
FB200_AMG_FUSED=0 timeout 300 python bench.py --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/c7_1gpu_unfused.json 2> gpurun_out/c7_1gpu_unfused.err; echo "1gpu unfused exit $?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c7_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", {a: d["ssh"][a] for a in ("ms_per_step", "iterations", "ms_per_iteration")}, "parity", d.get("parity"))
        print("   ", {n: (round(v["ms"] / v["launches"], 4), v["launches"]) for n, v in k.items() if n.startswith("k_amg") or n.startswith("k_spmv") or n.startswith("nccl") or n.startswith("k_flat") or n in ("k_x_update","k_apply_scalars","halo_exchange_flat","k_halo_pack","k_halo_unpack")})
    except Exception as e:
        print(f, "unreadable:", e)
PY
for f in gpurun_out/c7_*.err; do echo $f; tail -n 3 $f; done
