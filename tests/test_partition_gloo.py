"""CPU test of the multi-process path (world_size 2, gloo): rank 0 partitions the global model and ships each rank its
partition (partition.scatter_local_models); every rank then drives the oracle on its partition with point-to-point
halo exchanges at the reference's exchange points and must reproduce the undivided oracle bit for bit."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from harness import partition as P
    from oracle import binding as ob
    from tests import util
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        hm = util.build_model("band")                      # every rank builds it only to check the result
        part = P.contiguous_partition(hm, world)
        lm = P.scatter_local_models(hm if rank == 0 else None, part, dist)
        nl, ntr = lm.ipar("nl"), lm.ipar("num_tracers")
        N = lm.ipar("myDim_nod2D") + lm.ipar("eDim_nod2D")
        Eall = lm.ipar("myDim_elem2D") + lm.ipar("eDim_elem2D") + lm.ipar("eXDim_elem2D")
        o = ob.TracerOracle(lm)
        for tr in range(1, ntr + 1):
            o.phase(1, tr)
            P.exchange_dist(lm, o.work("tr_xy", (2, nl - 1, Eall)), "elem", dist)
            o.phase(2, tr)
            o.phase(3, tr)
            P.exchange_dist(lm, o.work("fct_LO", (nl - 1, N)), "nod", dist)
            o.phase(4, tr)
            P.exchange_dist(lm, o.work("fct_plus", (nl - 1, N)), "nod", dist)
            P.exchange_dist(lm, o.work("fct_minus", (nl - 1, N)), "nod", dist)
            o.phase(5, tr)
            P.exchange_dist(lm, lm["tr_arr"][tr - 1], "nod", dist)
        o.final()
        ob.TracerOracle(hm).step()
        ids = lm["myList_nod2D"] - 1
        ok = bool(np.array_equal(lm["tr_arr"], hm["tr_arr"][:, ids]))
        q.put((rank, ok, int(lm.ipar("myDim_nod2D"))))
    finally:
        dist.destroy_process_group()


def test_two_process_gloo_partitioned_oracle():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29650 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(60)
    assert len(res) == 2 and all(ok for _, ok, _ in res), res
    assert sum(n for _, _, n in res) > 0
