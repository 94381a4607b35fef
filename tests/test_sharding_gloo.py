"""N>1 host logic on CPU: two processes, gloo backend.  The per-shard engine here is the numpy restatement
(test infrastructure); on a GPU box the same ShardedCluster drives gsx.Snapshot engines over NCCL."""
import importlib
import os
import socket
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if "gsx" not in sys.modules:  # spawned workers import this module without conftest
    sys.modules["gsx"] = importlib.import_module("gpushare-scheduler-extender_b200")

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gsx import sharding, synth
from oracle import soa_reference as S

WORLD = 2


class NumpyEngine:
    """Shard engine for the CPU test: flat-array restatement of Filter, Bind and pod events on the shard's own
    nodes.  Same interface as sharding.SnapshotEngine: tensors in, tensors out, on the collective's device."""

    def __init__(self, sc, sh, rank):
        lo, hi = sh.node_range(rank)
        self.n = hi - lo
        self.cap, self.ndev, self.cls = sc.dev_capacity()[lo:hi], sc.dev_count()[lo:hi], sc.node_class()[lo:hi]
        recs, pos = sh.route(sc.gs_pod_records(), rank)
        recs["flags"][~S.legal(recs, self.ndev)] = 0  # as gs_pods_load stores them (nodeinfo.go:98-108)
        self.table = np.zeros(sc.n_pods + 4096, synth.POD_DT)
        self.table[pos] = recs
        self.used = S.used_from_records(recs, self.n, sc.max_devs, self.ndev)

    def _keys(self):
        return S.node_keys(self.cap, self.used, self.ndev, self.cls)

    def filter(self, req, stride_words):
        key, elig = self._keys()
        bm = S.pack_bitmap(S.filter_pass(req.numpy().view(np.uint64), key, elig), stride_words)
        return torch.from_numpy(bm.view(np.int32))

    def filter_compact(self, req, rank_stride):
        key, elig = self._keys()
        r = req.numpy().view(np.uint64)
        srt = np.sort(r)
        rank = np.zeros(rank_stride, np.uint16)
        rank[: self.n] = np.where(elig, np.searchsorted(srt, key, side="right"), 0)
        cutoff = (1 + np.searchsorted(srt, r, side="left")).astype(np.uint16)
        return torch.from_numpy(rank.view(np.int16)), torch.from_numpy(cutoff.view(np.int16))

    def bind(self, binds):
        return S.bind_sequential(binds, self.cap, self.used, self.ndev, self.cls, self.table)

    def apply_events(self, events):
        S.apply_events(events, self.table, self.used, self.ndev)


def _scenario(sc):
    reqs = synth.make_requests(300, "adversarial", 5, max_req=20000).astype(np.uint64)
    rng = np.random.default_rng(3)
    nodes = rng.integers(0, sc.n_nodes, 400)
    binds = synth.bind_records(nodes, rng.choice(synth.CATALOGUE[:6], 400), first_slot=sc.n_pods)
    ev = np.concatenate([synth.add_events(sc.n_pods + 1000 + np.arange(200), rng.integers(0, sc.n_nodes, 200),
                                          rng.integers(0, 4, 200), rng.choice(synth.CATALOGUE[:5], 200)),
                         synth.remove_events(rng.integers(0, sc.n_pods, 300))])
    return reqs, binds, ev


def _worker(rank, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        sc = synth.make_cluster(777, 4, 16384, running=4000, seed=21, quirks=True, fill=0.8)
        sh = sharding.NodeSharding(sc.n_nodes, WORLD)
        cl = sharding.ShardedCluster(sh, rank, NumpyEngine(sc, sh, rank), dist)
        reqs, binds, ev = _scenario(sc)
        bm = cl.filter(reqs if rank == 0 else None)
        devs = cl.bind(binds if rank == 0 else None)
        bm2 = cl.filter(reqs if rank == 0 else None)  # after the binds
        cl.apply_events(ev if rank == 0 else None)
        req_t = cl.broadcast_requests(reqs if rank == 0 else None)
        g = cl.filter_device(req_t)                    # gathered shard-major form, after the events
        rk, cut = cl.filter_compact_device(req_t)
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), bm=bm, devs=devs, bm2=bm2, g=g.numpy().view(np.uint32),
                 rk=rk.numpy().view(np.uint16), cut=cut.numpy().view(np.uint16))
    finally:
        dist.destroy_process_group()


def test_two_rank_sharded_filter_and_bind_match_unsharded(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(port, str(tmp_path)), nprocs=WORLD, join=True)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    for k in ("bm", "devs", "bm2", "g", "rk", "cut"):
        assert (r0[k] == r1[k]).all()                               # every rank ends with the same result
    # unsharded reference
    sc = synth.make_cluster(777, 4, 16384, running=4000, seed=21, quirks=True, fill=0.8)
    reqs, binds, ev = _scenario(sc)
    cap, ndev, cls = sc.dev_capacity(), sc.dev_count(), sc.node_class()
    recs = sc.gs_pod_records()
    used = S.used_from_records(recs, sc.n_nodes, sc.max_devs, ndev)
    sh = sharding.NodeSharding(sc.n_nodes, WORLD)
    W = sh.world * sh.words_per_shard
    key, elig = S.node_keys(cap, used, ndev, cls)

    def globalise(passes):  # shard-major column order: shard g's nodes start at word g * words_per_shard
        out = np.zeros((passes.shape[0], W * 32), dtype=bool)
        for g in range(sh.world):
            lo, hi = sh.node_range(g)
            out[:, g * sh.per_shard: g * sh.per_shard + (hi - lo)] = passes[:, lo:hi]
        return S.pack_bitmap(out, W)

    assert (r0["bm"] == globalise(S.filter_pass(reqs, key, elig))).all()
    table = np.zeros(sc.n_pods + 4096, synth.POD_DT)
    table[: sc.n_pods] = recs
    want = S.bind_sequential(binds, cap, used, ndev, cls, table)
    assert (r0["devs"] == want).all()
    key, elig = S.node_keys(cap, used, ndev, cls)
    assert (r0["bm2"] == globalise(S.filter_pass(reqs, key, elig))).all()
    # the events, then the gathered forms: shard-major bitmap and compact ranks / cutoffs
    table["flags"][: sc.n_pods][~S.legal(recs, ndev)] = 0
    S.apply_events(ev, table, used, ndev)
    key, elig = S.node_keys(cap, used, ndev, cls)
    want3 = S.filter_pass(reqs, key, elig)
    assert (sh.assemble_bitmap(r0["g"]) == globalise(want3)).all()
    rank = sh.assemble_rank(r0["rk"])
    assert ((rank[None, :] >= r0["cut"][:, None]) == want3).all()


def test_sharding_geometry():
    sh = sharding.NodeSharding(50_000, 8)
    assert sh.per_shard == 6272 and sh.words_per_shard == 196
    assert sh.node_range(7) == (43904, 50000)
    assert sharding.NodeSharding(10, 4).per_shard == 32
    ev = np.concatenate([synth.add_events([1, 2], [5, 40], [0, 1], [7, 9]), synth.remove_events([3])])
    out, pos = sharding.NodeSharding(64, 2).route_events(ev, 1)
    assert pos.tolist() == [1, 2] and out["node_idx"].tolist() == [8, 0]
