set -x
python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c1_pytest.log
tail -5 gpurun_out/c1_pytest.log
python bench.py --nodes 1000000 --steps 5 --warmup 3 --no-cpu-baseline --pc amg > gpurun_out/c1_b1m_amg.json 2> gpurun_out/c1_b1m_amg.err; echo "exit $?"
python bench.py --nodes 1000000 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --pc bj > gpurun_out/c1_b1m_bj.json 2> gpurun_out/c1_b1m_bj.err; echo "exit $?"
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --pc amg > gpurun_out/c1_b5m_amg.json 2> gpurun_out/c1_b5m_amg.err; echo "exit $?"
tail -c 600 gpurun_out/c1_b1m_amg.err gpurun_out/c1_b5m_amg.err
