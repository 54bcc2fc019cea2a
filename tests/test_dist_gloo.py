"""world_size-2 gloo tests (CPU) of the N>1 path: the C++ drivers shard every
batch of tuples over ranks and meet in one all-reduce(min) per PC round / one
all-gather per lattice level.  The evaluator is the checker (oracle), injected
through the test seam; the collectives are real torch.distributed ones."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    import openbnsl_b200
    from helpers import OracleEvaluator, load_golden, pdag_to_adj
    from openbnsl_b200.dist import TorchComm
    L = openbnsl_b200.openbnsllib
    comm = TorchComm()
    out = {}
    g, net, codes, cards = load_golden("alarm_3k_s1")
    ods = oracle.Dataset(codes, cards)
    for e in g["pc"]:
        if not e["stable"]:
            continue
        T = L.citest.ChiSquare(e["level"]) if e["kind"] == 0 else L.citest.GSquare(e["level"])
        ev = OracleEvaluator(ods)
        got = L._testing.pc(net.n, ev, T, e["max_cond_vars"], True, comm, 0)
        out[("pc", e["kind"])] = (pdag_to_adj(got, net.n).tolist() == e["adj"], ev.n_ci,
                                  L.structure_learning.last_pc_stats()["total_tests"])
    g, net, codes, cards = load_golden("asia_10k_s0")
    ods = oracle.Dataset(codes, cards)
    for e in g["exhaustive_search"]:
        ev = OracleEvaluator(ods)
        got = L._testing.exhaustive_search(net.n, ev, L.score.BDeu(e["ess"]), e["max_parents"], 2, comm)
        out[("es", e["max_parents"])] = (pdag_to_adj(got, net.n).tolist() == e["adj"], ev.n_sc,
                                         L.structure_learning.last_es_stats()["scores"])
    # simulated annealing: chains are dealt to ranks (rank r runs chains r, r + world, ...); one
    # all-gather of (best score, adjacency) picks the winner, which must be the reference's graph
    import json
    with open(os.path.join(ROOT, "tests", "golden", "simulated_annealing.json")) as f:
        sa = json.load(f)["asia_10k_s0"]
    sa_out = []
    for e in sa:
        ev = OracleEvaluator(ods)
        got = L._testing.simulated_annealing(net.n, ev, L.score.BDeu(e["ess"]), e["max_parents"], e["max_iters"],
                                             e["init_temp"], e["cooling_rate"], e["is_deterministic"], e["seed"],
                                             e["num_chains"], comm)
        sa_out.append((pdag_to_adj(got, net.n).tolist() == e["adj"], ev.n_sc, e["num_chains"]))
    out["sa"] = sa_out
    out["collectives"] = comm.n_collectives
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_world2_sharded_drivers():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=500) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    keys = [k for k in res[0] if k not in ("collectives", "sa")]
    assert keys
    for k in keys:
        ok0, mine0, total0 = res[0][k]
        ok1, mine1, total1 = res[1][k]
        assert ok0 and ok1, k                         # both ranks learn the reference's graph
        assert total0 == total1
        assert mine0 + mine1 == total0, k             # the work was split, not replicated
        assert abs(mine0 - mine1) <= max(16, total0 // 4), k
    assert res[0]["collectives"] == res[1]["collectives"] > 0
    for (ok0, n0, chains), (ok1, n1, _) in zip(res[0]["sa"], res[1]["sa"]):
        assert ok0 and ok1                            # every rank returns the reference's DAG
        assert n0 > 0 and n1 > 0                      # and both scored their own chains
