"""Two-GPU NCCL test of the sharded drivers (skipped on a one-GPU box).

Each rank owns one GPU, the dataset is replicated, every PC round / lattice
level is sharded over the ranks (shard_min_bytes=0 forces sharding even for
small rounds), the collectives run inside libopenbnsl_b200 (obn_comm_init) and the learned graphs must equal the golden ones produced by the
unmodified reference."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, json
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
os.environ["OBN_DEVICE"] = str(local)
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import openbnsl_b200
from openbnsl_b200.dist import TorchComm, init_native_comm
from helpers import load_golden, pdag_to_adj
L = openbnsl_b200.openbnsllib
L.profile_enable(True)
comm = init_native_comm(L)            # NCCL owned by the library: the product path
assert comm.world == 2 and comm.rank == rank
host_comm = TorchComm()               # host-side collectives through torch.distributed: the A/B
ok = True
g, net, codes, cards = load_golden("alarm_3k_s1")
df = L.base.DataframeWrapper.from_codes(codes, net.names, [int(c) for c in cards])
for e in g["pc"]:
    if not e["stable"]:
        continue
    T = L.citest.ChiSquare(e["level"]) if e["kind"] == 0 else L.citest.GSquare(e["level"])
    for cm, overlap in ((comm, True), (comm, False), (host_comm, False)):
        got = L.structure_learning.pc(df, T, e["max_cond_vars"], True, comm=cm, shard_min_bytes=0, overlap=overlap)
        ok &= pdag_to_adj(got, net.n).tolist() == e["adj"]
n_pc_collectives = L.profile_read(False)["collectives"]
ok &= n_pc_collectives > 0
g, net, codes, cards = load_golden("asia_10k_s0")
df = L.base.DataframeWrapper.from_codes(codes, net.names, [int(c) for c in cards])
for e in g["exhaustive_search"]:
    got = L.structure_learning.exhaustive_search(df, L.score.BDeu(e["ess"]), e["max_parents"], comm=comm, shard_min_bytes=0)
    ok &= pdag_to_adj(got, net.n).tolist() == e["adj"]
# fast mode: marginalisation lattice, blocks split over the ranks, F all-gathered in place over NCCL
from openbnsl_b200 import networks as nw
from helpers import ranked
net = nw.random_binary(12, seed=9)
codes, cards = ranked(net.sample(6000, 3))
df = L.base.DataframeWrapper.from_codes(codes, net.names, cards)
a = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), 11, exact_order=0, comm=comm, lattice=True, dp_shard_min=1000000000)
sa = L.structure_learning.last_es_stats()
b = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), 11, exact_order=0, lattice=True)
sb = L.structure_learning.last_es_stats()
ok &= sa["lattice"] and sb["lattice"] and sa["best_score"] == sb["best_score"]
ok &= pdag_to_adj(a, net.n).tolist() == pdag_to_adj(b, net.n).tolist()
# the same with the DP sharded by child from the very first level (every rank keeps its children's
# columns only; (sum, child) merged by one all-gather per level), and from a middle level on
for dp_min in (1, 200):
    before = L.profile_read(False)["collectives"]
    a2 = L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), 11, exact_order=0, comm=comm, lattice=True, dp_shard_min=dp_min)
    s2 = L.structure_learning.last_es_stats()
    ok &= s2["lattice"] and s2["best_score"] == sb["best_score"]
    ok &= pdag_to_adj(a2, net.n).tolist() == pdag_to_adj(b, net.n).tolist()
    ok &= L.profile_read(False)["collectives"] - before > 8
t = torch.tensor([int(ok), int(L.profile_read(False)["collectives"])], device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print("MULTI_GPU_RESULT", json.dumps({"ok": bool(t[0].item()), "collectives": int(t[1].item())}))
L.comm_destroy()
dist.destroy_process_group()
'''


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.timeout(600)
def test_two_gpu_pc_and_es_match_golden(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    env = dict(os.environ)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script)],
                       capture_output=True, text=True, env=env, timeout=540)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-8000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("MULTI_GPU_RESULT")]
    assert line, r.stdout[-2000:]
    import json
    res = json.loads(line[0].split(" ", 1)[1])
    assert res["ok"] and res["collectives"] > 0
