# GPU call 10 (2 GPUs): BiCGStab iterations as the body of a conditional while-node; fused coarse levels with 1024-thread blocks
set -x
mkdir -p gpurun_out
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
B="bench.py --gpus 2 --nodes 1000000 --no-e2e --no-cpu-baseline --steps 5"
$T $B > gpurun_out/c10_1m_while.json 2> gpurun_out/c10_1m_while.err; echo "while exit $?"
FB200_SSH_WHILE=0 $T $B --no-parity > gpurun_out/c10_1m_poll8.json 2> gpurun_out/c10_1m_poll8.err; echo "poll8 exit $?"
FB200_AMG_FUSED=1 $T $B --no-parity > gpurun_out/c10_1m_while_fused.json 2> gpurun_out/c10_1m_while_fused.err; echo "fused exit $?"
FB200_AMG_FUSED=1 FB200_SSH_OVERLAP=4 $T $B --no-parity > gpurun_out/c10_1m_while_fused_ov4.json 2> gpurun_out/c10_1m_while_fused_ov4.err; echo "fused ov4 exit $?"
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_ssh.py -x -q > gpurun_out/c10_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c10_pytest.log
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/c10_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "tracer_ms", round(d["tracer_ms_per_step"], 2), "ssh", {a: d["ssh"][a] for a in ("ms_per_step", "iterations", "ms_per_iteration")}, "parity", (d.get("parity") or {}).get("ok"))
    except Exception as e:
        print(f, "unreadable:", e)
PY
for f in gpurun_out/c10_*.err; do echo $f; tail -n 3 $f; done
