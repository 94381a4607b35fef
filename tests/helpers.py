"""Shared test plumbing: builds the object-level oracle cluster that corresponds to a SynthCluster."""
from __future__ import annotations

import numpy as np

import gsx
from gsx import synth
from oracle import oracle as O
from oracle import soa_reference as S

PHASE = {synth.PENDING: O.PENDING, synth.RUNNING: O.RUNNING, synth.SUCCEEDED: O.SUCCEEDED, synth.FAILED: O.FAILED}


def oracle_cluster(sc: "synth.SynthCluster", build: bool = True) -> "O.Cluster":
    c = O.Cluster()
    for i in range(sc.n_nodes):
        if not sc.node_listed[i]:
            continue
        t, k = int(sc.node_total[i]), int(sc.node_count[i])
        c.set_node(synth.node_name(i), mem=None if t < 0 else t, count=None if k < 0 else k)
    for j in range(sc.n_pods):
        mem = int(sc.pod_mem[j])
        p = O.Pod(f"run-{j}", uid=f"uid-run-{j}", node=synth.node_name(int(sc.pod_node[j])), phase=PHASE[int(sc.pod_phase[j])],
                  annotations={O.IDX: int(sc.pod_gpu[j]), O.POD: mem}, limits=[mem])
        c.list_pod(p)
    if build:
        err = c.build_cache()
        # pods on unlisted nodes abort BuildCache in the reference (cache.go:65-68); the generator
        # only produces them when quirks are on, and then the oracle is fed pod by pod instead.
        assert err is None, err
    return c


def oracle_cluster_tolerant(sc: "synth.SynthCluster") -> "O.Cluster":
    """Like BuildCache but skipping pods whose node is unknown (what a per-pod syncPod would do)."""
    c = O.Cluster()
    for i in range(sc.n_nodes):
        if not sc.node_listed[i]:
            continue
        t, k = int(sc.node_total[i]), int(sc.node_count[i])
        c.set_node(synth.node_name(i), mem=None if t < 0 else t, count=None if k < 0 else k)
    for j in range(sc.n_pods):
        mem = int(sc.pod_mem[j])
        p = O.Pod(f"run-{j}", uid=f"uid-run-{j}", node=synth.node_name(int(sc.pod_node[j])), phase=PHASE[int(sc.pod_phase[j])],
                  annotations={O.IDX: int(sc.pod_gpu[j]), O.POD: mem}, limits=[mem])
        if p.annotation_mem <= 0:
            continue
        c.add_or_update_pod(p)
    return c


def all_names(sc) -> list:
    return [synth.node_name(i) for i in range(sc.n_nodes)]


def pending_pod(i: int, req: int) -> "O.Pod":
    return O.Pod(f"pend-{i}", uid=f"uid-pend-{i}", limits=[int(req)])


def soa_state(sc):
    """(dev_capacity, used, dev_count, node_class, records) as the numpy reference sees them."""
    recs = sc.gs_pod_records()
    used = S.used_from_records(recs, sc.n_nodes, sc.max_devs, sc.dev_count())
    return sc.dev_capacity(), used, sc.dev_count(), sc.node_class(), recs


def load_snapshot(sc, extra_pods: int = 0, device: int = 0) -> "gsx.Snapshot":
    snap = gsx.Snapshot(sc.n_nodes, sc.max_devs, max(1, sc.n_pods + extra_pods), device=device)
    snap.node_upsert(sc.gs_nodes())
    if sc.n_pods:
        snap.pods_load(sc.gs_pod_records())
    snap.rebuild()
    return snap
