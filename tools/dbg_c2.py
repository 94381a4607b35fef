import importlib, sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200"); sys.modules["gsx"] = gsx
from gsx import synth
from oracle import soa_reference as S
from tests import helpers as H
sc = synth.make_config("C2")
reqs = np.r_[synth.make_requests(500, "adversarial", 0xC2), synth.make_requests(500, "catalogue", 0xC2)].astype(np.uint64)
cap, used, ndev, cls, _ = H.soa_state(sc)
key, elig = S.node_keys(cap, used, ndev, cls)
want = S.filter_pass(reqs, key, elig)
with H.load_snapshot(sc) as s:
    got = gsx.unpack_bitmap(s.filter_batch(reqs), sc.n_nodes)
bad = np.argwhere(got != want)
print("mismatches", len(bad))
if len(bad):
    pods = np.unique(bad[:, 0]); nodes = np.unique(bad[:, 1])
    print("pods", len(pods), pods[:20], "reqs", reqs[pods[:20]])
    print("nodes", len(nodes), nodes[:20], "keys", key[nodes[:20]], "words", np.unique(nodes // 32)[:20])
    order = np.argsort(reqs >> np.uint64(9), kind="stable")
    pos = np.empty(1000, int); pos[order] = np.arange(1000)
    print("approx grouped positions of bad pods (tile = pos//64):", np.unique(pos[pods] // 64))
    for p, n in bad[:10]:
        print(f"pod {p} req {reqs[p]} node {n} key {key[n]} elig {elig[n]} got {got[p, n]} want {want[p, n]}")
