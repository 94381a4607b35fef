"""The product's own annotation / limit parsers (C++ host mirror, pkg/utils/pod.go) against the CPU oracle. No GPU."""
import numpy as np

from gsx import ext as X
from oracle import oracle as O

STRINGS = ["", "0", "7", "+7", "-7", "007", "-0", "+", "-", " 1", "1 ", "1_0", "0x10", "1.5", "abc", "1e3", "--1",
           "9223372036854775807", "9223372036854775808", "-9223372036854775808", "-9223372036854775809",
           "99999999999999999999999", "99999999999999999999999x", "18446744073709551616", "184320"]


def test_atoi_matches_oracle():
    for s in STRINGS:
        assert X.atoi(s) == O.atoi(s), s
    rng = np.random.default_rng(0)
    alphabet = list("0123456789+-_ x")
    for _ in range(3000):
        s = "".join(rng.choice(alphabet, rng.integers(0, 24)))
        assert X.atoi(s) == O.atoi(s), s


def test_pod_parsers_match_oracle():
    rng = np.random.default_rng(1)
    for _ in range(500):
        ann = {}
        if rng.random() < 0.8:
            ann[O.IDX] = str(rng.choice(STRINGS))
        if rng.random() < 0.8:
            ann[O.POD] = str(rng.choice(STRINGS))
        if rng.random() < 0.3:
            ann["other"] = "x"
        limits = [None if rng.random() < 0.2 else int(rng.integers(-5, 2**40)) for _ in range(rng.integers(0, 4))]
        a, b = X.Pod("p", annotations=ann, limits=limits), O.Pod("p", annotations=ann, limits=limits)
        assert (a.gpu_id, a.annotation_mem, a.resource_mem) == (b.gpu_id, b.annotation_mem, b.resource_mem), (ann, limits)


def test_atoi_against_go_strconv_vectors():
    """Both restatements of strconv.Atoi -- the oracle's and the product's -- against the rows of Go's own
    strconv/atoi_test.go (tests/golden/atoi_vectors.json says where each row comes from)."""
    import json
    import os

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "atoi_vectors.json")
    doc = json.load(open(path))
    n = 0
    for table in ("parseInt64Tests", "parseUint64Tests_as_Atoi", "documented_behaviour"):
        for row in doc[table]:
            want = (row["value"], row["err"])
            assert O.atoi(row["in"]) == want, ("oracle", table, row)
            assert X.atoi(row["in"]) == want, ("product", table, row)
            n += 1
    assert n >= 30
