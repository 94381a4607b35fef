#!/usr/bin/env python
"""torchrun --nproc-per-node G tools/sharded_check.py : ShardedCluster over NCCL with Snapshot engines on G GPUs,
checked on rank 0 against the unsharded numpy restatement (development / GPU-box check, not the judged bench)."""
import importlib, os, sys
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200"); sys.modules["gsx"] = gsx
from gsx import sharding, synth
from oracle import soa_reference as S

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
sc = synth.make_config("C3")
sh = sharding.NodeSharding(sc.n_nodes, world)
lo, hi = sh.node_range(rank)
snap = gsx.Snapshot(sh.per_shard, sc.max_devs, sc.n_pods + 8192, device=lr)
nodes = sc.gs_nodes()[lo:hi].copy(); nodes["node_idx"] -= lo
snap.node_upsert(nodes)
recs = sc.gs_pod_records()
mine, pos = sh.route(recs, rank)
table = np.zeros(sc.n_pods, synth.POD_DT); table[pos] = mine
snap.pods_load(table); snap.rebuild()
cl = sharding.ShardedCluster(sh, rank, sharding.SnapshotEngine(snap), dist, device=f"cuda:{lr}")
reqs = synth.make_requests(2000, "adversarial", 9).astype(np.uint64)
bm = cl.filter(reqs if rank == 0 else None)
rng = np.random.default_rng(1)
bn = rng.integers(0, sc.n_nodes, 3000)
binds = synth.bind_records(bn, rng.choice(synth.CATALOGUE, 3000), first_slot=sc.n_pods)
devs = cl.bind(binds if rank == 0 else None)
if rank == 0:
    cap, ndev, cls = sc.dev_capacity(), sc.dev_count(), sc.node_class()
    used = S.used_from_records(recs, sc.n_nodes, sc.max_devs, ndev)
    key, elig = S.node_keys(cap, used, ndev, cls)
    W = sh.world * sh.words_per_shard
    passes = S.filter_pass(reqs, key, elig)
    out = np.zeros((len(reqs), W * 32), dtype=bool)
    for g in range(world):
        a, b = sh.node_range(g)
        out[:, g * sh.per_shard: g * sh.per_shard + (b - a)] = passes[:, a:b]
    assert (bm == S.pack_bitmap(out, W)).all(), "sharded filter differs"
    t = np.zeros(sc.n_pods + 8192, synth.POD_DT); t[: sc.n_pods] = recs
    want = S.bind_sequential(binds, cap, used, ndev, cls, t)
    assert (devs == want).all(), "sharded bind differs"
    print(f"sharded check ok on {world} GPUs: filter bitmap and bind choices match the unsharded restatement")
snap.close(); dist.destroy_process_group()
