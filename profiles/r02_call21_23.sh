# GPU calls 21-23: numbering study and BASELINE config 4 on several GPUs
#   21 (1 GPU):  python bench.py --nodes 1000000 --rcm --no-e2e --no-cpu-baseline                      -> r02_c21_bench_1m_rcm.json
#   22 (2 GPUs): torchrun --nproc-per-node 2 bench.py --gpus 2 --config farclike --no-e2e --no-cpu-baseline -> r02_c22_farclike_2gpu.json
#   23 (8 GPUs): the same with 8 ranks and --no-parity.  FAILED before any kernel ran: include/fesom_b200.h had been edited after the
#                last local build, every rank found libfesom_b200.so stale and all eight rebuilt and relinked it at the same time; one
#                rank loaded a half-linked file (r02_c23_farclike_8gpu_build_race.err.txt).  Fixed afterwards: the builds take a file lock
#                and replace their outputs atomically (fesom2_b200/build.py _BuildLock; eight concurrent loaders of a stale library pass
#                here).  The round's GPU budget ended with this call, so no 8-GPU figure exists from this session.
