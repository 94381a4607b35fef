"""SURVEY.md Appendix A, item by item, against the CPU oracle.  These pin down the edge-case
semantics of the Go sources (cited per test) that the CUDA path must reproduce."""
from oracle import oracle as O

MAXI = 2**63 - 1


def running(name, node, idx, mem, phase=O.RUNNING, uid=None, limits=None):
    return O.Pod(name, uid=uid, node=node, phase=phase, annotations={O.IDX: idx, O.POD: mem},
                 limits=limits if limits is not None else [mem])


# ---- strconv.Atoi (Go stdlib) as used at pkg/utils/pod.go:51,98 ----
def test_atoi():
    assert O.atoi("0") == (0, 0) and O.atoi("+7") == (7, 0) and O.atoi("-12") == (-12, 0)
    assert O.atoi("007") == (7, 0)
    for bad in ["", "+", "-", " 1", "1 ", "1_0", "0x10", "1.5", "abc", "1e3", "--1"]:
        assert O.atoi(bad) == (0, 1), bad
    assert O.atoi("9223372036854775807") == (MAXI, 0)
    assert O.atoi("9223372036854775808") == (MAXI, 2)
    assert O.atoi("-9223372036854775808") == (-(2**63), 0)
    assert O.atoi("-9223372036854775809") == (-(2**63), 2)
    assert O.atoi("99999999999999999999999") == (MAXI, 2)
    # overflow is reported where it happens, before a later bad byte is seen
    assert O.atoi("99999999999999999999999x") == (MAXI, 2)
    assert O.atoi("18446744073709551616") == (MAXI, 2)


def test_annotation_parsers():
    # pod.go:45-60: missing / unparseable / overflow -> -1
    assert O.Pod("a").gpu_id == -1
    assert O.Pod("a", annotations={O.IDX: "x"}).gpu_id == -1
    assert O.Pod("a", annotations={O.IDX: "99999999999999999999"}).gpu_id == -1
    assert O.Pod("a", annotations={O.IDX: "-3"}).gpu_id == -3
    assert O.Pod("a", annotations={O.IDX: "2"}).gpu_id == 2
    # pod.go:94-113: the Atoi error is ignored
    assert O.Pod("a").annotation_mem == 0
    assert O.Pod("a", annotations={O.POD: "abc"}).annotation_mem == 0
    assert O.Pod("a", annotations={O.POD: "-5"}).annotation_mem == 0
    assert O.Pod("a", annotations={O.POD: "99999999999999999999"}).annotation_mem == MAXI
    assert O.Pod("a", annotations={O.POD: "-99999999999999999999"}).annotation_mem == 0
    assert O.Pod("a", annotations={O.POD: "4096"}).annotation_mem == 4096
    # pod.go:146-155: sum of container limits
    assert O.Pod("a", limits=[1, None, 2]).resource_mem == 3
    assert O.Pod("a", limits=[]).resource_mem == 0 and not O.Pod("a", limits=[]).sharing


# ---- A. node ingestion ----
def test_a1_a2_capacity_truncation_and_missing_keys():
    c = O.Cluster()
    c.set_node("n", mem=100, count=3)          # nodeinfo.go:34: 100/3 = 33 truncated
    assert c.inspect_node("n") == ([33, 33, 33], [0, 0, 0])
    c.set_node("m", mem=100)                   # no gpu-count -> no devices
    assert c.inspect_node("m") == ([], [])
    c.set_node("z", count=4)                   # no gpu-mem -> devices of capacity 0
    assert c.inspect_node("z") == ([0, 0, 0, 0], [0, 0, 0, 0])


def test_a3_node_info_frozen_except_flip():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    c.add_or_update_pod(running("p", "n", 0, 10))
    c.set_node("n", mem=128, count=4)          # capacity change is ignored (cache.go:141-160)
    assert c.inspect_node("n") == ([32, 32], [10, 0])
    # flip total<=0 -> >0 replaces the NodeInfo and drops its podMaps (cache.go:145-157)
    c.set_node("f", mem=0, count=2)
    c.add_or_update_pod(running("q", "f", 1, 5))
    assert c.inspect_node("f") == ([0, 0], [0, 5])
    c.set_node("f", mem=40, count=2)
    assert c.inspect_node("f") == ([20, 20], [0, 0])
    # flip count<=0 -> >0
    c.set_node("g", mem=40)
    assert c.inspect_node("g") == ([], [])
    c.set_node("g", mem=40, count=4)
    assert c.inspect_node("g") == ([10, 10, 10, 10], [0, 0, 0, 0])


def test_a4_lister_consulted_first():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    assert c.get_node_info("n") is None
    c.delete_node("n")
    assert c.get_node_info("n") == 'node "n" not found'
    code, err = c.predicate(O.Pod("p", limits=[1]), "n")
    assert code == O.NOT_FOUND and err == 'node "n" not found'


# ---- B. pod -> device membership ----
def test_b5_b6_add_rules():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    assert c.add_or_update_pod(running("nonode", "", 0, 8)) is None       # cache.go:92-95
    assert c.add_or_update_pod(running("ghost", "nope", 0, 8)) == 'node "nope" not found'
    for idx in ["-1", "2", "x", "99999999999999999999"]:                  # nodeinfo.go:98-108
        p = running("bad" + idx, "n", idx, 8)
        assert c.add_or_update_pod(p) is None
        assert not c.known_pod(f"uid-default-bad{idx}")
    assert c.inspect_node("n") == ([32, 32], [0, 0])
    ok = running("ok", "n", 1, 8)
    c.add_or_update_pod(ok)
    assert c.known_pod("uid-default-ok") and c.inspect_node("n")[1] == [0, 8]


def test_b7_same_uid_replaces_and_moved_uid_is_double_counted():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    c.add_or_update_pod(running("p", "n", 0, 8, uid="U"))
    c.add_or_update_pod(running("p", "n", 0, 12, uid="U"))      # deviceinfo.go:63 replace
    assert c.inspect_node("n")[1] == [12, 0]
    c.add_or_update_pod(running("p", "n", 1, 12, uid="U"))      # nodeinfo.go:93-105 never removes
    assert c.inspect_node("n")[1] == [12, 12]


def test_b8_remove_uses_id_of_passed_pod():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    c.add_or_update_pod(running("p", "n", 0, 8, uid="U"))
    c.remove_pod(running("p", "n", 1, 8, uid="U"))              # wrong device: nothing removed there
    assert c.inspect_node("n")[1] == [8, 0] and not c.known_pod("U")
    c.remove_pod(running("p", "n", 0, 8, uid="U"))
    assert c.inspect_node("n")[1] == [0, 0]


def test_b9_build_cache_admission_and_phase():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    c.list_pod(running("run", "n", 0, 8))
    c.list_pod(running("done", "n", 0, 4, phase=O.SUCCEEDED))   # admitted, counts 0 (deviceinfo.go:46-49)
    c.list_pod(running("fail", "n", 1, 4, phase=O.FAILED))
    c.list_pod(running("pend", "n", 1, 2, phase=O.PENDING))
    c.list_pod(O.Pod("noann", node="n", phase=O.RUNNING, limits=[8]))       # annotation mem 0 -> skipped
    c.list_pod(running("unsched", "", 0, 8))                                 # no node -> skipped
    assert c.build_cache() is None
    assert c.inspect_node("n")[1] == [8, 2]
    assert c.known_pod("uid-default-done") and not c.known_pod("uid-default-noann")


def test_build_cache_aborts_on_unknown_node():
    c = O.Cluster()
    c.list_pod(running("p", "missing", 0, 8))
    assert c.build_cache() == 'node "missing" not found'


# ---- C. used / available ----
def test_c10_c11_wraparound():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    c.add_or_update_pod(running("big", "n", 0, 40))             # oversubscribed: 32 - 40 wraps
    pod = O.Pod("q", limits=[10**15])
    assert c.predicate(pod, "n")[0] == O.PASS                   # nodeinfo.go:260 huge avail fits anything
    assert c.allocate_gpu_id("n", pod) == (0, True)
    # smaller request: device 1 (avail 32) is strictly smaller than the wrapped device 0
    assert c.allocate_gpu_id("n", O.Pod("r", limits=[16])) == (1, True)
    c.add_or_update_pod(O.Pod("of", node="n", phase=O.RUNNING, annotations={O.IDX: 1, O.POD: "9" * 30}))
    assert c.inspect_node("n")[1] == [40, MAXI]


# ---- D. filter ----
def test_d13_d14_d15_filter_rules():
    c = O.Cluster()
    c.set_node("share", mem=64, count=2)
    c.set_node("noshare", count=2)
    c.set_node("nodevs", mem=64)
    c.add_or_update_pod(running("a", "share", 0, 30))
    c.add_or_update_pod(running("b", "share", 1, 20))
    names = ["share", "noshare", "ghost", "nodevs", "share"]
    pod = O.Pod("p", limits=[12])
    assert list(c.filter_codes(pod, names)) == [O.PASS, O.NOT_SHARE, O.NOT_FOUND, O.INSUFFICIENT, O.PASS]
    assert c.predicate(pod, "noshare")[1] == "The node noshare is not for GPU share, need skip"
    assert list(c.filter_codes(O.Pod("p", limits=[13]), names))[0] == O.INSUFFICIENT
    # req == 0 passes wherever there is a device; a sharing node without devices is Insufficient
    zero = O.Pod("z", limits=[])
    assert list(c.filter_codes(zero, names)) == [O.PASS, O.NOT_SHARE, O.NOT_FOUND, O.INSUFFICIENT, O.PASS]
    body = c.filter_json(pod, names)
    assert body == ('{"Nodes":null,"NodeNames":["share","share"],"FailedNodes":{"ghost":"node \\"ghost\\" not found",'
                    '"nodevs":"Insufficient GPU Memory in one device",'
                    '"noshare":"The node noshare is not for GPU share, need skip"},"Error":""}')
    assert c.filter_json(pod, []) == '{"Nodes":null,"NodeNames":[],"FailedNodes":{},"Error":""}'


def test_negative_request_wraps_to_huge():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    neg = O.Pod("neg", limits=[-1])
    assert c.predicate(neg, "n")[0] == O.INSUFFICIENT           # uint(-1) > any avail
    c.add_or_update_pod(running("big", "n", 0, 33))             # avail wraps to 2^64-1
    assert c.predicate(neg, "n")[0] == O.PASS


# ---- E. bind ----
def test_e17_e18_zero_request_and_error_text():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    z = O.Pod("z", uid="uz", limits=[])
    c.list_pod(z)
    dev, err = c.bind("z", "default", "uz", "n")
    assert dev == -1 and err == "The node  can't place the pod z in ns default"   # nodeinfo.go:170
    big = O.Pod("big", uid="ub", limits=[33])
    c.list_pod(big)
    assert c.bind("big", "default", "ub", "n") == (-1, "The node  can't place the pod big in ns default")
    assert c.api_binds == 0 and c.inspect_node("n")[1] == [0, 0]


def test_e17_ties_pick_lowest_index_and_strict_less():
    c = O.Cluster()
    c.set_node("n", mem=400, count=4)
    for d, m in enumerate([10, 50, 50, 90]):
        c.add_or_update_pod(running(f"r{d}", "n", d, m))
    assert c.allocate_gpu_id("n", O.Pod("p", limits=[50])) == (1, True)   # avail 90,50,50,10 -> tie 1 vs 2
    assert c.allocate_gpu_id("n", O.Pod("p", limits=[10])) == (3, True)
    assert c.allocate_gpu_id("n", O.Pod("p", limits=[91])) == (-1, False)


def test_e19_e20_commit_and_failures():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    for i, mode in enumerate([O.API_OTHER_ERROR, O.API_CONFLICT_GET_FAILS, O.API_CONFLICT_TWICE, O.API_BIND_FAILS]):
        p = O.Pod(f"f{i}", uid=f"uf{i}", limits=[8])
        c.list_pod(p)
        dev, err = c.bind(f"f{i}", "default", f"uf{i}", "n", api_mode=mode)
        assert dev == 0 and err != ""
        assert c.inspect_node("n")[1] == [0, 0]                # no commit on any failure
    p = O.Pod("ok", uid="uok", limits=[8])
    c.list_pod(p)
    c.api_set_pod(p)
    assert c.bind("ok", "default", "uok", "n", api_mode=O.API_CONFLICT_THEN_OK) == (0, "")
    assert c.inspect_node("n")[1] == [8, 0]
    assert not c.known_pod("uok")                              # Allocate does not rememberPod
    # later informer update re-adds the same UID on the same device: idempotent (cache.go:102-104)
    bound = c.api_pod("default", "ok").set_node("n")
    c.add_or_update_pod(bound)
    assert c.inspect_node("n")[1] == [8, 0] and c.known_pod("uok")
    # binpack: next 8 goes to device 0 again (avail 24 < 32)
    q = O.Pod("q", uid="uq", limits=[8])
    c.list_pod(q)
    assert c.bind("q", "default", "uq", "n") == (0, "")
    assert c.inspect_node("n")[1] == [16, 0]


def test_bind_negative_request_commits_zero():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    c.add_or_update_pod(running("big", "n", 1, 33))            # device 1 wraps to 2^64-1: fits uint(-1)
    neg = O.Pod("neg", uid="un", limits=[-1])
    c.list_pod(neg)
    assert c.bind("neg", "default", "un", "n") == (1, "")
    assert c.inspect_node("n")[1] == [0, 33]                   # annotation "-1" counts as 0 (pod.go:99-101)


def test_e22_bind_ignores_sharing_flag_and_uid_check():
    c = O.Cluster()
    c.set_node("z", count=2)                                   # capacity-0 devices
    p = O.Pod("p", uid="u1", limits=[1])
    c.list_pod(p)
    assert c.bind("p", "default", "u1", "z") == (-1, "The node  can't place the pod p in ns default")
    assert c.bind("p", "default", "other", "z")[1] == 'pods "p" not found'
    c.api_set_pod(O.Pod("p", uid="u2", limits=[1]))
    assert c.bind("p", "default", "other", "z")[1] == \
        "The pod p in ns default's uid is u2, and it's not equal with expected other"
    assert c.bind("nope", "default", "x", "z")[1] == 'pod "nope" not found'
    assert c.bind("p", "default", "u1", "ghost")[1] == 'node "ghost" not found'


# ---- F. event gating ----
def test_f23_event_path():
    c = O.Cluster()
    c.set_node("n", mem=64, count=2)
    nogpu = O.Pod("nogpu", limits=[])
    c.list_pod(nogpu)
    c.informer_add(nogpu)
    assert c.queue_len == 0                                    # FilterFunc: IsGPUsharingPod
    p = O.Pod("p", uid="U", limits=[8])
    c.list_pod(p)
    c.informer_add(p)
    assert c.queue_len == 1 and c.drain() is None
    assert not c.known_pod("U")                                # no node yet -> no-op
    # bind writes annotations; informer update of an unknown pod with IDX >= 0 enqueues
    p2 = p.clone().set_annotation(O.IDX, 1).set_annotation(O.POD, 8).set_node("n")
    c.list_pod(p2)
    c.informer_update(p, p2)
    assert c.queue_len == 1 and c.drain() is None
    assert c.known_pod("U") and c.inspect_node("n")[1] == [0, 8]
    # update of a known, not complete pod: ignored
    p3 = p2.clone().set_phase(O.RUNNING)
    c.list_pod(p3)
    c.informer_update(p2, p3)
    assert c.queue_len == 0
    # known and complete: removed
    p4 = p3.clone().set_phase(O.SUCCEEDED)
    c.list_pod(p4)
    c.informer_update(p3, p4)
    assert c.queue_len == 1 and c.drain() is None
    assert not c.known_pod("U") and c.inspect_node("n")[1] == [0, 0]
    # delete path uses the stashed object
    q = running("q", "n", 0, 5, uid="Q")
    c.list_pod(q)
    c.informer_add(q)
    c.drain()
    assert c.inspect_node("n")[1] == [5, 0]
    c.unlist_pod("default", "q")
    c.informer_delete(q)
    assert c.drain() is None and c.inspect_node("n")[1] == [0, 0]
