# GPU call 12 (1 GPU): threads per row of the multigrid kernels, 500 k rows, graph on
mkdir -p gpurun_out
run() { echo "== $*"; env "$@" python profiles/ssh_timeline.py --nodes 500000 --solves 6 --graph 1 2>&1 | tail -2; }
run FB200_AMG_NQ=1
run FB200_AMG_NQ=8
run FB200_AMG_NQ=4
run FB200_AMG_NQ=2
run FB200_AMG_NQ=8 FB200_AMG_NQ0=2
run FB200_AMG_NQ=8 FB200_AMG_NQ0=4
run FB200_AMG_NQ=8 FB200_AMG_FUSED=1
run FB200_AMG_NQ=8 FB200_SSH_WHILE=0
