# full ncu capture of the column kernels (batched loads) at 1 M nodes; preceded by the same command without ncu
set -x
mkdir -p gpurun_out
B1="python bench.py --nodes 1000000 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
$B1 > gpurun_out/plain1.log 2>&1 && ncu --set full --clock-control none --import-source on -k 'regex:^k_(ab2_g|grad_elem2|edge_adv2|ho_edge_fb|lo2|fct_bounds2|edge_diff2|update2|impl_diff)' -c 13 -o gpurun_out/prof_r02_tracer $B1 > gpurun_out/ncu1.log 2>&1
tail -n 3 gpurun_out/ncu1.log
ls -la gpurun_out/*.ncu-rep
