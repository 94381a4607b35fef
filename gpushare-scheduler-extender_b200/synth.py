"""Seeded synthetic clusters for the configurations of BASELINE.json.

Pure numpy; emits (a) object-level facts -- node capacities, pod annotations / phases / limits -- that
tests feed to the CPU oracle, and (b) the flat struct-of-arrays records the C ABI takes.  Everything is
an integer number of MiB, as the reference's extended resources are.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from .capi import BIND_DT, EVENT_DT, NODE_DT, POD_DT, NODE_NOT_FOUND, NODE_NOT_SHARE, NODE_OK, POD_COUNTED, POD_VALID

PENDING, RUNNING, SUCCEEDED, FAILED = 0, 1, 2, 3

CATALOGUE = np.array([256, 512, 1024, 2048, 4096, 8192, 16384, 24576, 40960, 81920], dtype=np.int64)
_ZIPF_W = 1.0 / np.arange(1, len(CATALOGUE) + 1) ** 1.1
_ZIPF_W /= _ZIPF_W.sum()
HETERO_CAPS = np.array([16384, 24576, 32768, 40960, 81920, 143360, 184320, 196608], dtype=np.int64)

# name -> (nodes, devices per node, per-device capacity MiB or None for heterogeneous, running pods, pending pods)
CONFIGS = {
    "C1": dict(n_nodes=10, n_devs=4, dev_cap=16384, running=0, pending=100, seed=0xC1),
    "C2": dict(n_nodes=1_000, n_devs=8, dev_cap=184320, running=20_000, pending=1_000, seed=0xC2),
    "C3": dict(n_nodes=5_000, n_devs=8, dev_cap=None, running=100_000, pending=5_000, seed=0xC3),
    "C4": dict(n_nodes=50_000, n_devs=8, dev_cap=184320, running=1_000_000, pending=10_000, seed=0xC4),
    "C5": dict(n_nodes=50_000, n_devs=8, dev_cap=184320, running=1_000_000, pending=1_000, seed=0xC5),
}


def node_name(i: int) -> str:
    return "node-%06d" % i


@dataclass
class SynthCluster:
    n_nodes: int
    max_devs: int
    node_total: np.ndarray     # int64[N] Capacity["shared-gpu/gpu-mem"], -1 = key absent
    node_count: np.ndarray     # int64[N] Capacity["shared-gpu/gpu-count"], -1 = key absent
    node_listed: np.ndarray    # bool[N]  False = name absent from the node lister
    pod_node: np.ndarray       # int32[R] node index
    pod_gpu: np.ndarray        # int64[R] SHARED_GPU_MEM_IDX annotation value (may be -1 / out of range)
    pod_mem: np.ndarray        # int64[R] SHARED_GPU_MEM_POD annotation value
    pod_phase: np.ndarray      # uint8[R]
    meta: dict = field(default_factory=dict)

    @property
    def n_pods(self) -> int:
        return len(self.pod_node)

    # ---- what NewNodeInfo derives (pkg/cache/nodeinfo.go:29-45, pkg/utils/node.go:6-30) ----
    def dev_count(self) -> np.ndarray:
        return np.maximum(self.node_count, 0).astype(np.uint32)

    def dev_capacity(self) -> np.ndarray:
        total = np.maximum(self.node_total, 0)  # missing key -> 0; the generator never emits negatives
        cnt = np.maximum(self.node_count, 1)
        cap = np.where(self.node_count > 0, total // cnt, 0)
        return cap.astype(np.uint64)

    def node_class(self) -> np.ndarray:
        cls = np.full(self.n_nodes, NODE_OK, dtype=np.uint8)
        cls[np.maximum(self.node_total, 0) <= 0] = NODE_NOT_SHARE
        cls[~self.node_listed] = NODE_NOT_FOUND
        return cls

    def gs_nodes(self) -> np.ndarray:
        a = np.zeros(self.n_nodes, dtype=NODE_DT)
        a["node_idx"] = np.arange(self.n_nodes, dtype=np.uint32)
        a["dev_capacity"] = self.dev_capacity()
        a["dev_count"] = self.dev_count()
        a["node_class"] = self.node_class()
        return a

    def gs_pod_records(self) -> np.ndarray:
        """One record per running pod, slot = pod index.  The annotation id is clamped to int16 with
        -1 for anything that is not a device index (pkg/utils/pod.go:45-60 + nodeinfo.go:98-108)."""
        r = np.zeros(self.n_pods, dtype=POD_DT)
        gid = self.pod_gpu
        ok = (gid >= 0) & (gid < 32767)
        r["mem"] = np.maximum(self.pod_mem, 0).astype(np.uint64)
        r["node_idx"] = self.pod_node.astype(np.uint32)
        r["dev"] = np.where(ok, gid, -1).astype(np.int16)
        counted = (self.pod_phase != SUCCEEDED) & (self.pod_phase != FAILED)
        r["flags"] = (POD_VALID | np.where(counted, POD_COUNTED, 0)).astype(np.uint8)
        # pods on nodes absent from the lister can never be added (cache.go:97-100)
        r["flags"][~self.node_listed[self.pod_node]] = 0
        return r


def make_cluster(n_nodes: int, n_devs: int = 8, dev_cap: int | None = 184320, running: int = 0,
                 seed: int = 0, quirks: bool = True, fill: float | None = None) -> SynthCluster:
    """SURVEY.md 8(d) node and running-pod distributions.  `dev_cap=None` draws a per-node capacity
    from HETERO_CAPS (config 3).  `quirks=False` gives a plain cluster (every node listed, sharing,
    full device count)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    if dev_cap is None:
        cap = HETERO_CAPS[rng.integers(0, len(HETERO_CAPS), n_nodes)]
    else:
        cap = np.full(n_nodes, dev_cap, dtype=np.int64)
    count = np.full(n_nodes, n_devs, dtype=np.int64)
    listed = np.ones(n_nodes, dtype=bool)
    total_extra = np.zeros(n_nodes, dtype=np.int64)
    noshare = np.zeros(n_nodes, dtype=bool)
    if quirks:
        u = rng.random(n_nodes)
        noshare = u < 0.01                                   # 1 %: no gpu-mem capacity
        listed = ~((u >= 0.01) & (u < 0.015))                # 0.5 %: not in the node lister
        few = (u >= 0.015) & (u < 0.025)                     # 1 %: 1, 2 or 4 devices
        count[few] = rng.choice(np.array([1, 2, 4]), int(few.sum()))
        count = np.minimum(count, n_devs)
        ragged = (u >= 0.025) & (u < 0.03) & (count > 1)     # 0.5 %: total % count != 0
        total_extra[ragged] = rng.integers(1, count[ragged])
    total = cap * count + total_extra
    total[noshare] = -1

    # running pods
    R = running
    pod_node = rng.integers(0, n_nodes, R).astype(np.int32)
    pod_gpu = (rng.random(R) * count[pod_node]).astype(np.int64)
    pod_mem = CATALOGUE[rng.choice(len(CATALOGUE), R, p=_ZIPF_W)]
    if fill is not None and R:
        # scale pod sizes so devices end up about `fill` full on average (makes small clusters interesting)
        mean_pods = max(R / max(1, int(count.sum())), 1e-9)
        target = fill * float(np.mean(cap)) / mean_pods
        pod_mem = np.maximum(1, (pod_mem * (target / float((CATALOGUE * _ZIPF_W).sum()))).astype(np.int64))
    # reject pods that would overfill their device (in arrival order), re-draw up to 4 times
    keep = np.ones(R, dtype=bool)
    dev_cap_of = np.where(count > 0, np.maximum(total, 0) // np.maximum(count, 1), 0)
    for _ in range(4):
        if R == 0:
            break
        bin_id = pod_node.astype(np.int64) * n_devs + pod_gpu
        order = np.argsort(bin_id, kind="stable")
        sb = bin_id[order]
        sm = np.where(keep[order], pod_mem[order], 0)
        cs = np.cumsum(sm)
        seg_start = np.r_[True, sb[1:] != sb[:-1]]
        base = np.maximum.accumulate(np.where(seg_start, cs - sm, 0))
        within = cs - base
        over = within > dev_cap_of[pod_node[order]]
        bad = np.zeros(R, dtype=bool)
        bad[order] = over & keep[order]
        if not bad.any():
            break
        nb = int(bad.sum())
        pod_node[bad] = rng.integers(0, n_nodes, nb)
        pod_gpu[bad] = (rng.random(nb) * count[pod_node[bad]]).astype(np.int64)
        pod_mem[bad] = CATALOGUE[rng.choice(4, nb)]
    if quirks and R:
        v = rng.random(R)
        bad_id = v < 0.002                                   # 0.2 %: id out of range or -1
        pod_gpu[bad_id] = np.where(rng.random(int(bad_id.sum())) < 0.5, -1, count[pod_node[bad_id]] + 3)
        oversub = (v >= 0.002) & (v < 0.003)                 # 0.1 %: deliberate oversubscription -> uint wrap
        pod_mem[oversub] = dev_cap_of[pod_node[oversub]] + rng.integers(1, 4096, int(oversub.sum()))
    ph = rng.random(R)
    pod_phase = np.full(R, RUNNING, dtype=np.uint8)
    pod_phase[ph < 0.01] = PENDING
    pod_phase[(ph >= 0.01) & (ph < 0.02)] = SUCCEEDED
    pod_phase[(ph >= 0.02) & (ph < 0.03)] = FAILED
    return SynthCluster(n_nodes, n_devs, total, np.where(count >= 0, count, -1), listed, pod_node, pod_gpu,
                        pod_mem.astype(np.int64), pod_phase, meta=dict(seed=seed, quirks=quirks))


def make_config(name: str, **over) -> SynthCluster:
    c = dict(CONFIGS[name])
    c.update(over)
    return make_cluster(c["n_nodes"], c["n_devs"], c["dev_cap"], c["running"], c["seed"])


def make_requests(n: int, kind: str = "adversarial", seed: int = 0, max_req: int = 184320) -> np.ndarray:
    """Pending-pod requests (int64, the sum of container limits).  `catalogue`: the ten sizes, Zipf(1.1);
    `adversarial`: uniform integers in [1, max_req] (about n distinct rows).  0.1 % are zero."""
    rng = np.random.Generator(np.random.PCG64(seed ^ 0x5EED))
    if kind == "catalogue":
        req = CATALOGUE[rng.choice(len(CATALOGUE), n, p=_ZIPF_W)].astype(np.int64)
    elif kind == "adversarial":
        req = rng.integers(1, max_req + 1, n, dtype=np.int64)
    else:
        raise ValueError(kind)
    req[rng.random(n) < 0.001] = 0
    return req


def bind_records(node_idx, req, first_slot: int, counted=True) -> np.ndarray:
    """gs_bind_req array for binds of pods whose limits sum to `req` (int64)."""
    req = np.asarray(req, dtype=np.int64)
    b = np.zeros(len(req), dtype=BIND_DT)
    b["req"] = req.astype(np.uint64)                        # uint(int): negatives wrap (nodeinfo.go:217)
    b["commit_mem"] = np.maximum(req, 0).astype(np.uint64)  # what Atoi("%d" % req) reads back (pod.go:98-101)
    b["node_idx"] = np.asarray(node_idx, dtype=np.uint32)
    b["slot"] = first_slot + np.arange(len(req), dtype=np.uint32)
    b["counted"] = np.asarray(counted, dtype=np.uint8)
    return b


def add_events(slot, node_idx, dev, mem, counted=True) -> np.ndarray:
    e = np.zeros(len(np.atleast_1d(slot)), dtype=EVENT_DT)
    e["slot"], e["node_idx"], e["dev"], e["mem"] = slot, node_idx, dev, mem
    e["op"] = 0
    e["counted"] = np.asarray(counted, dtype=np.uint8)
    return e


def remove_events(slot) -> np.ndarray:
    e = np.zeros(len(np.atleast_1d(slot)), dtype=EVENT_DT)
    e["slot"] = slot
    e["op"] = 1
    return e
