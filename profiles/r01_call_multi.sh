# multi-GPU check (gpurun --gpus 4): the partitioned parity tests (2 and 4 ranks) and a strong-scaling bench line
set -x
nvidia-smi -L | head -8
python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/c2_pytest_multi.log 2>&1; echo "pytest exit $?" >> gpurun_out/c2_pytest_multi.log
tail -5 gpurun_out/c2_pytest_multi.log
N=$(nvidia-smi -L | wc -l)
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --nodes 2000000 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/c2_b2m_multi.json 2> gpurun_out/c2_b2m_multi.err; echo "exit $?"
tail -c 1500 gpurun_out/c2_b2m_multi.err
tail -c 2500 gpurun_out/c2_b2m_multi.json
