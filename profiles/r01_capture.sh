# Round-1 ncu evidence (run under gpurun from the repo root; outputs land in gpurun_out/ and are summarised into profiles/)
#   1. launch list of the default bench command (5 M nodes) with per-launch device time
#   2. full captures of every tracer kernel and of the SSH iteration kernels (BiCGStab + multigrid V-cycle) at 1 M nodes
# Each ncu run is preceded by the same command without ncu (must exit 0).
set -x
B5="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
B1="python bench.py --nodes 1000000 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
$B5 > gpurun_out/plain5.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_5m.csv $B5 > gpurun_out/ncu5.log 2>&1
$B1 > gpurun_out/plain1.log 2>&1 && ncu --set full --clock-control none --import-source on -k 'regex:^k_(ab2|ho_edge|ho_edge_fb|update|lo|fct_bounds|impl_diff|diff_edge|grad_elem|clamp_salt)' -c 11 -o gpurun_out/prof_r01_tracer $B1 > gpurun_out/ncu1.log 2>&1
# one BiCGStab iteration = spmv_alpha, enter + 12 cycle kernels, spmv_sigma, x_update, enter + 12 cycle kernels = 29 launches
ncu --set full --clock-control none --import-source on -k 'regex:^k_(spmv_alpha|spmv_sigma|x_update|amg_enter|amg_resid|amg_restrict|amg_prolong|amg_post)' -s 87 -c 29 -o gpurun_out/prof_r01_ssh $B1 > gpurun_out/ncu1b.log 2>&1
tail -n 2 gpurun_out/ncu5.log gpurun_out/ncu1.log gpurun_out/ncu1b.log
ls -la gpurun_out/
