# GPU call 9 (2 GPUs): restricted additive Schwarz overlap of the per-rank V-cycle, 0/1/2/4 layers, METIS partitions
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/c9_pytest_multi.log 2>&1; echo "pytest multi exit $?"; tail -3 gpurun_out/c9_pytest_multi.log
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
B="bench.py --gpus 2 --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5"
for ov in 2 0 1 4; do
FB200_SSH_OVERLAP=$ov $T $B $([ $ov != 2 ] && echo --no-parity) > gpurun_out/c9_1m_ov$ov.json 2> gpurun_out/c9_1m_ov$ov.err; echo "1m ov$ov exit $?"
done
FB200_SSH_COMM=nccl $T $B --no-parity > gpurun_out/c9_1m_ov2_nccl.json 2> gpurun_out/c9_1m_ov2_nccl.err; echo "nccl exit $?"
for ov in 2 4; do
FB200_SSH_OVERLAP=$ov $T bench.py --gpus 2 --no-cpu-baseline --steps 5 $([ $ov != 2 ] && echo "--no-parity --no-e2e") > gpurun_out/c9_5m_ov$ov.json 2> gpurun_out/c9_5m_ov$ov.err; echo "5m ov$ov exit $?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c9_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        k = d.get("kernels", {})
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", {a: d["ssh"][a] for a in ("ms_per_step", "iterations", "ms_per_iteration")}, "e2e", d.get("e2e", {}).get("ms_per_step"), "parity", (d.get("parity") or {}).get("ok"))
        print("   ", {n: (round(v["ms"] / v["launches"], 4), v["launches"]) for n, v in k.items() if n.startswith("k_amg") or n.startswith("k_spmv") or n.startswith("nccl") or n.startswith("k_flat") or n in ("k_x_update","k_apply_scalars","halo_exchange_flat","k_halo_pack","k_halo_unpack")})
    except Exception as e:
        print(f, "unreadable:", e)
PY
for f in gpurun_out/c9_*.err; do echo $f; tail -n 3 $f; done
