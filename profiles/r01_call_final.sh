# what the driver runs at round end: GPU tests, smoke, the default bench line, the reference arm
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/f_pytest.log; tail -4 gpurun_out/f_pytest.log
python __graft_entry__.py smoke > gpurun_out/f_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/f_smoke.log
python bench.py > gpurun_out/f_bench_5m.json 2> gpurun_out/f_bench_5m.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/f_bench_ref.json 2> gpurun_out/f_bench_ref.err; echo "ref exit $?"
tail -c 400 gpurun_out/f_bench_5m.err gpurun_out/f_bench_ref.err
