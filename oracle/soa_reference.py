"""numpy restatement of the hot path on flat arrays -- the large-size checker.

TEST INFRASTRUCTURE ONLY (same rules as oracle/gpushare_oracle.c).  The faithful C oracle works on
pod and node objects and is too slow beyond ~10^5 pods; this module restates the same decisions on the
struct-of-arrays inputs of the C ABI so that full-size configurations can be checked.  It is itself
checked against the C oracle in tests/test_soa_reference.py (CPU, every round).

All arithmetic is uint64 with wrap-around, as Go's `uint` (pkg/cache/nodeinfo.go:254-264).
"""
from __future__ import annotations

import numpy as np

NODE_OK, NODE_NOT_FOUND, NODE_NOT_SHARE = 0, 1, 2
POD_VALID, POD_COUNTED = 1, 2
PASS, NOT_FOUND, NOT_SHARE, INSUFFICIENT = 0, 1, 2, 3


def legal(recs, dev_count):
    """nodeinfo.go:98-108: the record's device must exist on its node."""
    node = recs["node_idx"].astype(np.int64)
    in_range = node < len(dev_count)
    nd = np.where(in_range, dev_count[np.minimum(node, len(dev_count) - 1)], 0)
    return in_range & (recs["dev"] >= 0) & (recs["dev"].astype(np.int64) < nd)


def used_from_records(recs, n_nodes: int, max_devs: int, dev_count) -> np.ndarray:
    """deviceinfo.go:41-54 summed per (node, device): valid, counted, legal records only."""
    used = np.zeros(n_nodes * max_devs, dtype=np.uint64)
    m = ((recs["flags"] & (POD_VALID | POD_COUNTED)) == (POD_VALID | POD_COUNTED)) & legal(recs, dev_count)
    bins = recs["node_idx"][m].astype(np.int64) * max_devs + recs["dev"][m].astype(np.int64)
    with np.errstate(over="ignore"):
        np.add.at(used, bins, recs["mem"][m])
    return used.reshape(n_nodes, max_devs)


def node_keys(dev_capacity, used, dev_count, node_class):
    """key[n] = max over the node's devices of (capacity - used) mod 2^64 (nodeinfo.go:254-264);
    eligible[n] = the predicate can pass at all (class OK and at least one device, nodeinfo.go:123)."""
    n, D = used.shape
    with np.errstate(over="ignore"):
        avail = dev_capacity.astype(np.uint64)[:, None] - used.astype(np.uint64)
    mask = np.arange(D)[None, :] < np.asarray(dev_count)[:, None]
    key = np.where(mask, avail, np.uint64(0)).max(axis=1) if D else np.zeros(n, np.uint64)
    eligible = (np.asarray(node_class) == NODE_OK) & (np.asarray(dev_count) > 0)
    return key.astype(np.uint64), eligible


def filter_pass(req_u64, key, eligible) -> np.ndarray:
    """[P][N] bool: Predicate.Handler over NodeInfo.Assume (predicate.go:15-39, nodeinfo.go:113-137)."""
    req = np.asarray(req_u64, dtype=np.uint64)
    return (key[None, :] >= req[:, None]) & eligible[None, :]


def filter_codes(req_u64, key, eligible, node_class) -> np.ndarray:
    p = filter_pass(req_u64, key, eligible)
    codes = np.where(p, PASS, INSUFFICIENT).astype(np.uint8)
    cls = np.asarray(node_class)
    codes[:, cls == NODE_NOT_SHARE] = NOT_SHARE
    codes[:, cls == NODE_NOT_FOUND] = NOT_FOUND
    return codes


def pack_bitmap(passes: np.ndarray, stride_words: int) -> np.ndarray:
    P, N = passes.shape
    bits = np.zeros((P, stride_words * 32), dtype=np.uint8)
    bits[:, :N] = passes
    return np.packbits(bits, axis=1, bitorder="little").view(np.uint32).reshape(P, stride_words)


def bind_sequential(binds, dev_capacity, used, dev_count, node_class, table=None):
    """allocateGPUID + commit in queue order (nodeinfo.go:139-252).  Mutates `used` (and `table`, the
    pod-record array, when given).  Returns chosen devices (-1 none, -2 node not found, -3 the bind's slot
    holds an entry of another node: an ABI rule of the device path, no counterpart in the reference, where a
    committed pod is a fresh map entry)."""
    out = np.empty(len(binds), dtype=np.int32)
    M = (1 << 64) - 1
    for i in range(len(binds)):
        b = binds[i]
        n = int(b["node_idx"])
        if n >= len(dev_count) or node_class[n] == NODE_NOT_FOUND:
            out[i] = -2
            continue
        req = int(b["req"])
        cand, cand_mem = -1, 0
        if req > 0:
            cap = int(dev_capacity[n])
            for d in range(int(dev_count[n])):
                avail = (cap - int(used[n, d])) & M
                if avail >= req and (cand == -1 or cand_mem > avail):
                    cand, cand_mem = d, avail
        if cand >= 0 and table is not None:
            old = table[int(b["slot"])]
            if (old["flags"] & POD_VALID) and int(old["node_idx"]) != n:
                cand = -3
        out[i] = cand
        if cand >= 0:
            if table is not None:
                old = table[int(b["slot"])]
                if (old["flags"] & 3) == 3:
                    used[int(old["node_idx"]), int(old["dev"])] = (int(used[int(old["node_idx"]), int(old["dev"])]) - int(old["mem"])) & M
                table[int(b["slot"])] = (int(b["commit_mem"]), n, cand, POD_VALID | (POD_COUNTED if b["counted"] else 0), 0)
            if b["counted"]:
                used[n, cand] = (int(used[n, cand]) + int(b["commit_mem"])) & M
    return out


def apply_events(events, table, used, dev_count):
    """AddOrUpdatePod / RemovePod on slots, strictly in order (cache.go:89-127).  Mutates table, used."""
    M = (1 << 64) - 1
    for e in events:
        slot = int(e["slot"])
        if slot >= len(table):
            continue
        if e["op"] > 1:
            continue
        if e["op"] == 0:
            n, d = int(e["node_idx"]), int(e["dev"])
            if n >= len(dev_count) or d < 0 or d >= int(dev_count[n]):
                continue
        old = table[slot]
        if (old["flags"] & 3) == 3:
            on, od = int(old["node_idx"]), int(old["dev"])
            used[on, od] = (int(used[on, od]) - int(old["mem"])) & M
        if e["op"] == 0:
            table[slot] = (int(e["mem"]), n, d, POD_VALID | (POD_COUNTED if e["counted"] else 0), 0)
            if e["counted"]:
                used[n, d] = (int(used[n, d]) + int(e["mem"])) & M
        else:
            table[slot] = (0, 0, 0, 0, 0)
