"""Node-axis sharding of the snapshot over the GPUs of one box (one process per GPU, torch.distributed).

Every (pod, node) Filter evaluation reads one node's row only, a Bind mutates one node's row only and a pod
event touches one node (pkg/cache/nodeinfo.go:113-137,209-252), so the node axis shards with no cross-node term:
rank g owns the contiguous node range [g * per, (g + 1) * per) with `per` a multiple of 32, so that shard bitmaps
are word aligned.  Pod records, events and binds are routed to the owner of their node (indices rebased); the
pending-pod requests are broadcast; filter bitmaps are all-gathered as [G][P][words_per_shard] and viewed as
global rows.  The per-shard compute is an "engine" -- in the product a gsx.Snapshot on the rank's GPU.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class NodeSharding:
    n_nodes: int
    world: int

    @property
    def per_shard(self) -> int:
        per = -(-self.n_nodes // self.world)
        return -(-per // 32) * 32

    @property
    def words_per_shard(self) -> int:
        return self.per_shard // 32

    def owner(self, node_idx):
        return np.asarray(node_idx, dtype=np.int64) // self.per_shard

    def local(self, node_idx):
        return np.asarray(node_idx, dtype=np.int64) % self.per_shard

    def node_range(self, rank: int):
        lo = rank * self.per_shard
        return lo, min(self.n_nodes, lo + self.per_shard)

    # ---- routing (host side, numpy structured arrays of the C-ABI record types) ----
    def route(self, recs: np.ndarray, rank: int, field: str = "node_idx"):
        """Records owned by `rank`, node indices rebased to the shard, plus their positions in `recs`."""
        mine = np.flatnonzero(self.owner(recs[field]) == rank)
        out = recs[mine].copy()
        out[field] = self.local(out[field])
        return out, mine

    def route_events(self, events: np.ndarray, rank: int):
        """Pod events: ADDs go to the owner of their node; REMOVEs name only a slot, so they go to every
        rank (a REMOVE of a slot the rank never filled is a no-op there)."""
        is_remove = events["op"] == 1
        mine = np.flatnonzero(is_remove | (self.owner(events["node_idx"]) == rank))
        out = events[mine].copy()
        out["node_idx"] = np.where(out["op"] == 1, 0, self.local(out["node_idx"]))
        return out, mine

    def assemble_bitmap(self, gathered: np.ndarray) -> np.ndarray:
        """[G][P][words_per_shard] shard-major bitmaps -> [P][G * words_per_shard] global rows."""
        g, p, w = gathered.shape
        assert g == self.world and w >= self.words_per_shard
        return np.ascontiguousarray(np.transpose(gathered[:, :, : self.words_per_shard], (1, 0, 2))).reshape(p, g * self.words_per_shard)


class ShardedCluster:
    """One rank's view: its NodeSharding, its engine, and the collectives around them."""

    def __init__(self, sharding: NodeSharding, rank: int, engine, dist=None, device="cpu"):
        """`device`: where collective buffers live -- "cpu" for gloo, the rank's cuda device for NCCL."""
        self.sh, self.rank, self.engine, self.dist, self.device = sharding, rank, engine, dist, device

    def filter(self, req_on_rank0):
        """Rank 0 supplies the requests; every rank returns the global [P][words] bitmap."""
        import torch

        dist = self.dist
        dev = self.device
        n = torch.tensor([0 if req_on_rank0 is None else len(req_on_rank0)], dtype=torch.int64, device=dev)
        if dist is not None:
            dist.broadcast(n, src=0)
        req = torch.zeros(int(n.item()), dtype=torch.int64, device=dev)
        if self.rank == 0:
            req.copy_(torch.from_numpy(np.ascontiguousarray(req_on_rank0, dtype=np.uint64).view(np.int64)))
        if dist is not None:
            dist.broadcast(req, src=0)
        local = np.ascontiguousarray(self.engine.filter(req.cpu().numpy().view(np.uint64), self.sh.words_per_shard))
        mine = torch.from_numpy(local.view(np.int32)).to(dev)
        if dist is not None:
            parts = [torch.empty_like(mine) for _ in range(self.sh.world)]
            dist.all_gather(parts, mine)
            gathered = np.stack([p.cpu().numpy().view(np.uint32) for p in parts])
        else:
            gathered = local[None]
        return self.sh.assemble_bitmap(gathered)

    def bind(self, binds_on_rank0):
        """Binds are applied by the owner of their node, in queue order per node; results return in queue order."""
        import torch

        dist = self.dist
        obj = [binds_on_rank0]
        if dist is not None:
            dist.broadcast_object_list(obj, src=0, device=torch.device(self.device))
        binds = obj[0]
        mine, pos = self.sh.route(binds, self.rank)
        res = np.full(len(binds), np.iinfo(np.int32).min, dtype=np.int32)
        if len(mine):
            res[pos] = self.engine.bind(mine)
        t = torch.from_numpy(res).to(self.device)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)  # exactly one rank holds a real value per bind
        return t.cpu().numpy()


class SnapshotEngine:
    """The product engine: a gsx.Snapshot on this rank's GPU (raises without CUDA: there is no CPU path)."""

    def __init__(self, snapshot):
        self.snap = snapshot

    def filter(self, req_u64, stride_words):
        return self.snap.filter_batch(req_u64, row_stride_words=stride_words)

    def bind(self, binds):
        return self.snap.bind_batch(binds)
