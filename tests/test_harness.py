"""CPU tests of the test / bench infrastructure itself: bench.py's workload options, the synthetic mesh numberings, the build lock."""
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def test_bench_configs_and_numberings(monkeypatch):
    import bench
    for cfg, nodes, nl in (("sphere5m", 5_000_000, 49), ("core2like", 127_000, 48), ("farclike", 640_000, 48)):
        monkeypatch.setattr(sys, "argv", ["bench.py", "--config", cfg])
        a = bench.parse()
        assert (a.nodes, a.nl) == (nodes, nl)
    monkeypatch.setattr(sys, "argv", ["bench.py", "--config", "farclike", "--nodes", "12000", "--nl", "9"])
    a = bench.parse()
    assert (a.nodes, a.nl) == (12000, 9)
    sizes = {}
    for name, kw in (("row-major", {}), ("random", {"shuffle": True}), ("rcm", {"rcm": True}), ("farclike", {"refine_north": 3.0})):
        bench.MESH_KW.clear()
        bench.MESH_KW.update(kw)
        hm, info = bench.build_workload(12000, 9)
        n = info["nod2D"]
        assert 0.8 * 12000 < n < 1.25 * 12000, (name, n)
        tri = hm["elem2D_nodes"].astype(np.int64)
        sizes[name] = float(np.abs(tri[:, 0] - tri[:, 1]).mean())       # index distance of two nodes of an element: the locality of the numbering
    bench.MESH_KW.clear()
    assert sizes["random"] > 10 * sizes["row-major"]
    assert sizes["rcm"] < 3 * sizes["row-major"]                        # the offline renumbering INTEGRATION.md recommends restores it


def _hold_lock(d, t, q):
    sys.path.insert(0, ROOT)
    from fesom2_b200.build import _BuildLock
    with _BuildLock(d):
        q.put(("in", time.time()))
        time.sleep(t)
        q.put(("out", time.time()))


def test_build_lock_serialises_builders(tmp_path):
    """several ranks that find the same library stale must not compile and link into the same files at once"""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_hold_lock, args=(str(tmp_path), 0.4, q)) for _ in range(3)]
    for p in ps:
        p.start()
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    ev = sorted((q.get(timeout=5) for _ in range(6)), key=lambda e: e[1])
    assert [e[0] for e in ev] == ["in", "out"] * 3                      # never two holders at the same time
