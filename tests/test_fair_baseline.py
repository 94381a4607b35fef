"""The "fair" CPU baseline that bench.py times (oracle/fair_baseline.c) takes the same decisions as the restatement."""
import numpy as np

from gsx import synth
from oracle import fair as F
from oracle import soa_reference as S
from tests import helpers as H


def test_fair_baseline_matches_restatement():
    rng = np.random.default_rng(1)
    for seed, (n, d) in enumerate([(777, 4), (2000, 8), (33, 3)]):
        sc = synth.make_cluster(n, d, 184320 if d == 8 else 16384, running=5 * n, seed=seed, quirks=True, fill=1.1)
        cap, used, ndev, cls, _ = H.soa_state(sc)
        key, elig = S.node_keys(cap, used, ndev, cls)
        reqs = rng.integers(0, 200000, 700).astype(np.uint64)
        reqs[::9] = 0
        reqs[1::13] = np.uint64(2**64 - 1) - rng.integers(0, 9, len(reqs[1::13])).astype(np.uint64)
        reqs[2::17] = np.uint64(2**32 - 1) + rng.integers(0, 3, len(reqs[2::17])).astype(np.uint64) - np.uint64(1)
        stride = (n + 31) // 32 + seed
        got = F.filter_bitmap(cap, used, ndev, cls, reqs, stride)
        assert (got == S.pack_bitmap(S.filter_pass(reqs, key, elig), stride)).all(), seed
    assert F.threads() >= 1
