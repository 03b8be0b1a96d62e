"""TEST INFRASTRUCTURE — memory-safety sweep of the kernels on the CPU: the parity tests of the GPU suite against the
emulated library built with -fsanitize=address (every emulated device allocation sits between heap redzones, so an
out-of-bounds index in a kernel is reported).  Run as
    FB200_EMU_ASAN=1 LD_PRELOAD=$(g++ -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0 python tests/emu/asan_sweep.py
(compute-sanitizer is not available on the GPU pool).  Result of the last run: profiles/r01_emulation_asan_sweep.txt"""
import sys, time, traceback
sys.path.insert(0, "/root/repo")
from tests import emu
from fesom2_b200 import api
api._lib = emu.load()
import tests.test_gpu_tracer as T, tests.test_gpu_ssh as S, tests.test_gpu_zz_zonal as z, tests.test_gpu_zz_means as m, tests.test_gpu_zz_ssh_assembly as a
jobs = []
def J(nm, f): jobs.append((nm, f))
for name, kw, opts in [("pi", {}, {}), ("soufflet", {}, {}), ("band", {}, {}), ("pi", {"w_split": True, "w_max_cfl": 0.2}, {}),
                       ("pi", {}, {"tra_adv_hor": "MUSCL"}), ("pi", {}, {"tra_adv_hor": "UPW1", "tra_adv_ver": "UPW1"}),
                       ("band", {}, {"tra_adv_ver": "CDIFF"}), ("band", {}, {"tra_adv_ver": "PPM"}), ("pi", {}, {"tra_adv_lim": "NONE", "tra_adv_hor": "MUSCL"}),
                       ("pi", {}, {"i_vert_diff": False})]:
    def f(name=name, kw=kw, opts=opts):
        hm, eng, o, g_tr, g_tro = T.run_pair(name, opts, kw, steps=1)
        eng.close()
        import numpy as np
        from tests import util
        assert util.rel_err(g_tr, hm["tr_arr"]) < 1e-12
    J("tracer %s %s %s" % (name, kw, opts), f)
J("cavity", lambda: T.test_cavity_patch_matches_oracle({}))
J("ale zstar", lambda: T.test_device_ale_producers_match_host_mirror("band", {"which_ale": "zstar", "water_flux": 2e-7}))
J("ale zlevel", lambda: T.test_device_ale_producers_match_host_mirror("band", {"which_ale": "zlevel", "water_flux": 2e-7}))
J("ale wsplit", lambda: T.test_device_ale_producers_match_host_mirror("pi", {"which_ale": "zstar", "w_split": True, "w_max_cfl": 0.2, "umax": 0.5}))
J("zonal", lambda: z.test_relax_zonal_temp_matches_oracle(True))
J("means", lambda: m.test_device_means_match_stream_restatement(4))
J("assembly", lambda: a.test_device_ssh_assembly_matches_host_mirror("pi", {"which_ale": "zstar"}))
J("psolve bj", lambda: S.test_psolve_matches_oracle("pi", 4))
J("psolve amg", lambda: S.test_psolve_multigrid_matches_reference_and_oracle("pi"))
J("psolve refresh", lambda: S.test_value_refresh_and_warm_start())
bad = 0
for nm, f in jobs:
    t = time.time()
    try:
        f(); print("ASAN-SWEEP", nm, "ok", round(time.time() - t, 1), flush=True)
    except Exception:
        bad += 1; print("ASAN-SWEEP", nm, "FAIL", flush=True); traceback.print_exc(limit=4)
print("ASAN-SWEEP done, failures:", bad)
