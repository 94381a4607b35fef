#!/usr/bin/env python
"""Development aid (GS_EXPAND_TRACE build): timeline of the k_prep / k_expand launches of a run of pipelined filter
calls: per launch, when its first / last block entered, passed its start wait, and left (globaltimer, us)."""
import ctypes as C, importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules["gsx"] = gsx
from gsx import capi, synth
cfg = synth.CONFIGS["C4"]
sc = synth.make_cluster(cfg["n_nodes"], cfg["n_devs"], cfg["dev_cap"], cfg["running"], seed=cfg["seed"])
P, N = cfg["pending"], sc.n_nodes
words = (N + 31) // 32
stride = (words + 7) // 8 * 8
snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + P)
snap.node_upsert(sc.gs_nodes()); snap.pods_load(sc.gs_pod_records()); snap.rebuild()
snap.set_option(1, int(os.environ.get("PIPE", "1")))
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(device=dev); torch.cuda.set_stream(st)
outs = [torch.empty((P, stride), dtype=torch.int32, device=dev) for _ in range(4)]
req = synth.make_requests(P, os.environ.get("KIND", "catalogue"), cfg["seed"])
d_req = torch.from_numpy(np.asarray(req, dtype=np.int64)).to(dev)
lib = capi.load()
for i in range(10): snap.filter_batch_device(d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
torch.cuda.synchronize()
assert lib.gs_debug_ktrace_reset() == 0
nsteps = 24
for i in range(nsteps): snap.filter_batch_device(d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
torch.cuda.synchronize()
lo = np.zeros((64, 2, 6), dtype=np.uint64); hi = np.zeros((64, 2, 6), dtype=np.uint64)
assert lib.gs_debug_ktrace_read(lo.ctypes.data_as(C.POINTER(C.c_uint64)), hi.ctypes.data_as(C.POINTER(C.c_uint64))) == 0
valid = [q for q in range(64) if hi[q, 1, 0] != 0]
order = sorted(valid, key=lambda q: int(lo[q, 1, 0]))
t0 = int(lo[order[0], 0, 0])
f = lambda v: f"{(int(v) - t0) / 1e3:7.1f}"
print("launch | prep: entry first..last, [atomic back first..last, counts complete first..last], past-wait last, exit last | expand: entry first..last, past-wait first..last, exit first..last")
for q in order[4:16]:
    print(f"{q:4d}   | {f(lo[q,0,0])}..{f(hi[q,0,0])} [{f(lo[q,0,3])}..{f(hi[q,0,3])} {f(lo[q,0,4])}..{f(hi[q,0,4])}] {f(hi[q,0,1])} {f(hi[q,0,2])} | {f(lo[q,1,0])}..{f(hi[q,1,0])}  {f(lo[q,1,1])}..{f(hi[q,1,1])}  {f(lo[q,1,2])}..{f(hi[q,1,2])}")

# per-tile phase stamps of the LAST launch (SM cycles, thread 0 of each CTA): medians over the CTAs' clean tiles
kT, kS = 6, 16
tr = np.zeros(148 * 5 * kT * kS, dtype=np.uint64)
assert lib.gs_debug_expand_trace(tr.ctypes.data_as(C.POINTER(C.c_uint64)), tr.size) == 0
tr = tr.reshape(148 * 5, kT, kS).astype(np.int64)
names = ["wait inputs 0-1", "phase A 1-2", "keys/ticket 2-3", "stores issued 3-6", "end sync 6-7", "to next tile 7-0'"]
for it in range(4):
    ok = (tr[:, it, 7] > tr[:, it, 0]) & (tr[:, it, 0] > 0) & ((tr[:, it, 8] & 0xffff) == 0)
    if ok.sum() < 10:
        continue
    t = tr[ok, it]
    d = [t[:, 1] - t[:, 0], t[:, 2] - t[:, 1], t[:, 3] - t[:, 2], t[:, 6] - t[:, 3], t[:, 7] - t[:, 6]]
    nxt = tr[ok, it + 1, 0] - t[:, 7]
    nxt_ok = tr[ok, it + 1, 7] > tr[ok, it + 1, 0]
    print(f"tile #{it} of a CTA, {int(ok.sum())} clean tiles: " + ", ".join(f"{n} {int(np.median(x))}" for n, x in zip(names, d)) +
          f", {names[5]} {int(np.median(nxt[nxt_ok])) if nxt_ok.any() else -1}; whole {int(np.median(t[:, 7] - t[:, 0]))} cycles")

# dense tiles of the last launch (thread 0's stamps)
dn = ["inputs 0-1", "phase A 1-2", "sort rows 2-3", "clean lanes' stores 3-4", "barrier 4-10", "keys land 10-11",
      "searches (warp 0) 11-12", "other warps 12-13", "ticket, request 13-5", "partial lanes' stores 5-6", "end 6-7"]
sl = [(0, 1), (1, 2), (2, 3), (3, 4), (4, 10), (10, 11), (11, 12), (12, 13), (13, 5), (5, 6), (6, 7)]
rows = []
for it in range(kT):
    ok = (tr[:, it, 7] > tr[:, it, 0]) & (tr[:, it, 0] > 0) & (((tr[:, it, 8] >> 16) & 1) == 1)
    if ok.any():
        rows.append(tr[ok, it])
if rows:
    t = np.concatenate(rows)
    print(f"{len(t)} dense tiles, partial words median {int(np.median(t[:, 8] & 0xffff))}: " +
          ", ".join(f"{n} {int(np.median(t[:, b] - t[:, a]))}" for n, (a, b) in zip(dn, sl)) +
          f"; whole {int(np.median(t[:, 7] - t[:, 0]))} cycles (max {int((t[:, 7] - t[:, 0]).max())})")

# sparse tiles of the last launch
sn = ["inputs 0-1", "phase A 1-2", "keys requested, ticket 2-3", "clean lanes' stores 3-4", "compares 4-5", "partial lanes' stores 5-6", "end 6-7"]
ss = [(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (5, 6), (6, 7)]
rows = []
for it in range(kT):
    ok = (tr[:, it, 7] > tr[:, it, 0]) & (tr[:, it, 0] > 0) & (((tr[:, it, 8] >> 16) & 1) == 0) & ((tr[:, it, 8] & 0xffff) != 0)
    if ok.any():
        rows.append(tr[ok, it])
if rows:
    t = np.concatenate(rows)
    print(f"{len(t)} sparse tiles, partial words median {int(np.median(t[:, 8] & 0xffff))}: " +
          ", ".join(f"{n} {int(np.median(t[:, b] - t[:, a]))}" for n, (a, b) in zip(sn, ss)) +
          f"; whole {int(np.median(t[:, 7] - t[:, 0]))} cycles (max {int((t[:, 7] - t[:, 0]).max())})")

# per-CTA view of the LAST launch: when each CTA classified its tiles (globaltimer, stamp 9), of what kind they were
gt = tr[:, :, 9]
kind = np.where(((tr[:, :, 8] >> 16) & 1) == 1, 2, np.where((tr[:, :, 8] & 0xffff) != 0, 1, 0))   # 0 clean, 1 sparse, 2 dense
valid = (tr[:, :, 7] > tr[:, :, 0]) & (tr[:, :, 0] > 0) & (gt > 0)
n_t = valid.sum(axis=1)
act = n_t > 0
if act.any():
    t0l = gt[valid].min()
    last_start = np.array([gt[c, :][valid[c]].max() if n_t[c] else 0 for c in range(gt.shape[0])])
    cyc_last = np.array([int(tr[c, n_t[c] - 1, 7] - tr[c, n_t[c] - 1, 0]) if 0 < n_t[c] <= kT else 0 for c in range(gt.shape[0])])
    end_est = (last_start - t0l) / 1e3 + cyc_last / 1965.0     # us after the first classification of the launch
    print(f"last launch: {int(act.sum())} CTAs; tiles per CTA (<= {kT} recorded): " + ", ".join(f"{k}: {int((n_t[act] == k).sum())}" for k in range(1, kT + 1)))
    print("end of a CTA's last recorded tile, us after the launch's first classification: min %.1f median %.1f p90 %.1f max %.1f" %
          (end_est[act].min(), np.median(end_est[act]), np.percentile(end_est[act], 90), end_est[act].max()))
    order_c = np.argsort(-end_est)
    for c in order_c[:8]:
        if not act[c]:
            continue
        ks = "".join("CSD"[kind[c, i]] for i in range(n_t[c]))
        st = ", ".join(f"{(gt[c, i] - t0l) / 1e3:.1f}" for i in range(n_t[c]))
        print(f"  CTA {c}: tiles {ks}, classified at {st} us, ends ~{end_est[c]:.1f} us")
    for k in (0, 1, 2):
        first = kind[:, 0] == k
        if (first & act).any():
            print(f"  CTAs whose first tile is {'clean sparse dense'.split()[k]}: {int((first & act).sum())}, end median {np.median(end_est[first & act]):.1f} max {end_est[first & act].max():.1f} us")
