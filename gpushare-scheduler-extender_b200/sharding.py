"""Node-axis sharding of the snapshot over the GPUs of one box (one process per GPU, torch.distributed).

Every (pod, node) Filter evaluation reads one node's row only, a Bind mutates one node's row only and a pod
event touches one node (pkg/cache/nodeinfo.go:113-137,209-252), so the node axis shards with no cross-node term:
rank g owns the contiguous node range [g * per, (g + 1) * per) with `per` a multiple of 32, so that shard bitmaps
are word aligned.  Pod records, events and binds are routed to the owner of their node (indices rebased); the
pending-pod requests are broadcast; filter results are gathered device to device, shard-major:

    bitmap form    [G][P][words_per_shard]   one all-gather of each rank's [P][words_per_shard] block
    compact form   rank [G][per_shard] u16 + cutoff [P] u16 (the cutoffs depend on the requests only and are the
                   same on every rank)

Nothing on the Filter path goes through host memory: the per-shard compute writes a device buffer and the
collective reads that buffer.  The per-shard compute is an "engine": in the product a gsx.Snapshot on the rank's
GPU (SnapshotEngine); the CPU tests of the routing and the collectives plug in a stand-in.
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

    @property
    def stride_words(self) -> int:
        """Row stride of a shard bitmap: whole 16-byte columns, so the bitmap kernel takes its vector path."""
        return -(-self.words_per_shard // 4) * 4

    def owner(self, node_idx):
        return np.asarray(node_idx, dtype=np.int64) // self.per_shard

    def local(self, node_idx):
        return np.asarray(node_idx, dtype=np.int64) % self.per_shard

    def node_range(self, rank: int):
        lo = min(self.n_nodes, rank * self.per_shard)
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
        """[G][P][>= words_per_shard] shard-major bitmaps -> [P][G * words_per_shard] global rows (host side, for
        a consumer that wants whole rows; the gathered form can equally be indexed by (shard, pod) directly)."""
        g, p, w = gathered.shape
        assert g == self.world and w >= self.words_per_shard
        return np.ascontiguousarray(np.transpose(gathered[:, :, : self.words_per_shard], (1, 0, 2))).reshape(p, g * self.words_per_shard)

    def assemble_rank(self, gathered_rank: np.ndarray) -> np.ndarray:
        """[G][per_shard] shard-major ranks -> [n_nodes]."""
        return np.ascontiguousarray(gathered_rank).reshape(-1)[: self.n_nodes]


class ShardedCluster:
    """One rank's view: its NodeSharding, its engine, and the collectives around them.  `device` is where the
    collective buffers live: the rank's cuda device for NCCL, "cpu" for the gloo tests."""

    def __init__(self, sharding: NodeSharding, rank: int, engine, dist=None, device="cpu"):
        self.sh, self.rank, self.engine, self.dist, self.device = sharding, rank, engine, dist, device
        self._gathered = None
        self._gathered_rank = None
        self._peer = None  # (symmetric-memory handle, [two gathered rank buffers], their peer pointers, step counter)

    # ---- small host arrays from rank 0 to everybody (requests, binds, events), as one device tensor ----
    def _bcast_array(self, arr_on_rank0, dtype: np.dtype) -> np.ndarray:
        import torch

        dtype = np.dtype(dtype)
        dist = self.dist
        if dist is None:
            return np.ascontiguousarray(arr_on_rank0, dtype=dtype)
        n = torch.tensor([0 if arr_on_rank0 is None else len(arr_on_rank0)], dtype=torch.int64, device=self.device)
        dist.broadcast(n, src=0)
        buf = torch.zeros(int(n.item()) * dtype.itemsize, dtype=torch.uint8, device=self.device)
        if self.rank == 0 and len(buf):
            buf.copy_(torch.from_numpy(np.ascontiguousarray(arr_on_rank0, dtype=dtype).view(np.uint8).reshape(-1)))
        if len(buf):
            dist.broadcast(buf, src=0)
        return buf.cpu().numpy().view(dtype)

    def broadcast_requests(self, req_on_rank0):
        """The pending pods' requests as an int64 tensor on `device`, identical on every rank."""
        import torch

        dist = self.dist
        n = torch.tensor([0 if req_on_rank0 is None else len(req_on_rank0)], dtype=torch.int64, device=self.device)
        if dist is not None:
            dist.broadcast(n, src=0)
        req = torch.zeros(int(n.item()), dtype=torch.int64, device=self.device)
        if self.rank == 0 and len(req):
            req.copy_(torch.from_numpy(np.ascontiguousarray(req_on_rank0, dtype=np.uint64).view(np.int64)))
        if dist is not None and len(req):
            dist.broadcast(req, src=0)
        return req

    # ---- Filter ----
    def filter_device(self, req):
        """`req`: int64 tensor on `device`, the same on every rank.  Returns the gathered shard-major bitmap
        [G][P][stride_words] (int32 tensor on `device`, valid until the next call)."""
        import torch

        local = self.engine.filter(req, self.sh.stride_words)  # [P][stride] int32 on `device`
        if self.dist is None:
            return local[None]
        shape = (self.sh.world,) + tuple(local.shape)
        if self._gathered is None or tuple(self._gathered.shape) != shape:
            self._gathered = torch.empty(shape, dtype=local.dtype, device=local.device)
        # (the concatenated view of the output is the form every backend accepts)
        self.dist.all_gather_into_tensor(self._gathered.view(-1, local.shape[-1]), local)
        return self._gathered

    def filter(self, req_on_rank0):
        """Rank 0 supplies the requests; every rank returns the global [P][words] bitmap on the host."""
        g = self.filter_device(self.broadcast_requests(req_on_rank0))
        return self.sh.assemble_bitmap(g.cpu().numpy().view(np.uint32))

    def enable_peer_gather(self):
        """Let the rank kernel itself do the all-gather of the compact Filter: the gathered [G][per_shard] array lives in
        symmetric memory (every GPU's copy mapped into every process over NVLink) and each rank's kernel stores its
        slice into all G copies; one device-side barrier per call replaces the NCCL collective.  Two buffers alternate,
        so a rank that is one call ahead never writes what a peer may still be reading.  NCCL ranks on CUDA only."""
        import torch
        import torch.distributed._symmetric_memory as symm

        n = self.sh.world * self.sh.per_shard
        bufs, ptrs = [], []
        hdl = None
        for _ in range(2):
            t = symm.empty(n, dtype=torch.int16, device=self.device)
            h = symm.rendezvous(t, self.dist.group.WORLD)
            t.zero_()
            bufs.append(t)
            ptrs.append([int(p) for p in h.buffer_ptrs])
            hdl = h
        torch.cuda.synchronize()
        self.dist.barrier()
        self._peer = [hdl, bufs, ptrs, 0]

    def filter_compact_device(self, req):
        """Returns (rank [G][per_shard] int16 tensor, cutoff [P] int16 tensor) on `device`:
        pass(p, n) <=> rank[n // per_shard, n % per_shard] >= cutoff[p] (as unsigned 16-bit values).  Ranks are
        relative to the requests of the call, which are the same on every rank, and so are the cutoffs: only
        the ranks are gathered."""
        import torch

        if self._peer is not None:
            hdl, bufs, ptrs, step = self._peer
            k = step & 1
            self._peer[3] = step + 1
            cutoff = self.engine.filter_compact_gather(req, ptrs[k], self.rank * self.sh.per_shard, self.sh.per_shard)
            hdl.barrier(channel=0)  # every rank's slice has landed in every copy
            return bufs[k].view(self.sh.world, self.sh.per_shard), cutoff
        rank_local, cutoff = self.engine.filter_compact(req, self.sh.per_shard)
        if self.dist is None:
            return rank_local[None], cutoff
        shape = (self.sh.world,) + tuple(rank_local.shape)
        if self._gathered_rank is None or tuple(self._gathered_rank.shape) != shape:
            self._gathered_rank = torch.empty(shape, dtype=rank_local.dtype, device=rank_local.device)
        # (as bytes: not every backend gathers 16-bit integers)
        self.dist.all_gather_into_tensor(self._gathered_rank.view(-1).view(torch.uint8), rank_local.view(torch.uint8))
        return self._gathered_rank, cutoff

    # ---- Bind / events ----
    def bind(self, binds_on_rank0):
        """Binds are applied by the owner of their node, in queue order per node; results return in queue order
        on every rank (one small reduction: exactly one rank holds a real value per bind)."""
        import torch
        from .capi import BIND_DT

        binds = self._bcast_array(binds_on_rank0, BIND_DT)
        mine, pos = self.sh.route(binds, self.rank)
        res = np.full(len(binds), np.iinfo(np.int32).min, dtype=np.int32)
        if len(mine):
            res[pos] = self.engine.bind(mine)
        t = torch.from_numpy(res).to(self.device)
        if self.dist is not None and len(binds):
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.cpu().numpy()

    def apply_events(self, events_on_rank0):
        """Pod events go to the owner of their node (REMOVEs to everybody); no collective besides the broadcast."""
        from .capi import EVENT_DT

        events = self._bcast_array(events_on_rank0, EVENT_DT)
        mine, _ = self.sh.route_events(events, self.rank)
        if len(mine):
            self.engine.apply_events(mine)


def _current_stream_handle(device) -> int:
    """torch's current stream as a handle the C ABI takes.  The library reads a NULL stream as "the snapshot's own
    stream", so torch's default stream (handle 0) is named explicitly: cudaStreamLegacy."""
    import torch

    return torch.cuda.current_stream(device).cuda_stream or 1


class SnapshotEngine:
    """The product engine: a gsx.Snapshot on this rank's GPU (raises without CUDA: there is no CPU path).
    Filters read and write device memory only, on torch's current stream."""

    def __init__(self, snapshot):
        self.snap = snapshot
        self._out = None
        self._rank = None
        self._cutoff = None

    def filter(self, req, stride_words):
        import torch

        P = int(req.numel())
        if self._out is None or tuple(self._out.shape) != (P, stride_words):
            self._out = torch.zeros((P, stride_words), dtype=torch.int32, device=req.device)
        if P:
            self.snap.filter_batch_device(req.data_ptr(), P, self._out.data_ptr(), stride_words, _current_stream_handle(req.device))
        return self._out

    def filter_compact(self, req, rank_stride):
        import torch
        from .capi import MAX_BATCH

        P = int(req.numel())
        assert P <= MAX_BATCH, "one chunk of ranks per call"
        if self._rank is None or self._rank.numel() != rank_stride or self._cutoff is None or self._cutoff.numel() != P:
            self._rank = torch.zeros(rank_stride, dtype=torch.int16, device=req.device)
            self._cutoff = torch.zeros(P, dtype=torch.int16, device=req.device)
        if P:
            self.snap.filter_batch_compact_device(req.data_ptr(), P, self._rank.data_ptr(), rank_stride, self._cutoff.data_ptr(),
                                                  _current_stream_handle(req.device))
        return self._rank, self._cutoff

    def filter_compact_gather(self, req, dst_ptrs, dst_offset, slice_count):
        """The shard's ranks stored by the kernel into every array of dst_ptrs (this GPU's and the peers'); returns the
        cutoffs (the same on every rank)."""
        import torch

        P = int(req.numel())
        if self._cutoff is None or self._cutoff.numel() != P:
            self._cutoff = torch.zeros(P, dtype=torch.int16, device=req.device)
        self.snap.filter_batch_compact_device_gather(req.data_ptr(), P, dst_ptrs, dst_offset, slice_count, self._cutoff.data_ptr(),
                                                     _current_stream_handle(req.device))
        return self._cutoff

    def bind(self, binds):
        return self.snap.bind_batch(binds)

    def apply_events(self, events):
        self.snap.pods_apply(events)


def load_shard(sharding: NodeSharding, rank: int, sc, extra_pods: int = 0, device: int = 0):
    """A gsx.Snapshot holding rank's node range of the synthetic cluster `sc` (nodes and pod records routed and
    rebased, slots kept global so that events and binds name the same slots on every rank)."""
    from .capi import Snapshot

    lo, hi = sharding.node_range(rank)
    snap = Snapshot(max(1, sharding.per_shard), sc.max_devs, sc.n_pods + extra_pods + 1, device=device)
    nodes = sc.gs_nodes()[lo:hi].copy()
    nodes["node_idx"] -= lo
    if len(nodes):
        snap.node_upsert(nodes)
    recs = sc.gs_pod_records()
    foreign = sharding.owner(recs["node_idx"]) != rank
    recs["node_idx"] = sharding.local(recs["node_idx"])
    recs["flags"][foreign] = 0  # the slot exists on every rank; only the owner holds the entry
    snap.pods_load(recs)
    snap.rebuild()
    return snap
