import sys, numpy as np
raw = open(sys.argv[1], "rb").read()
grid = int(np.frombuffer(raw[:4], np.uint32)[0])
a = np.frombuffer(raw[4:], np.uint64).reshape(grid, 4, 8).astype(np.int64)
# slots: 0 globaltimer start(ns), 1 clk start, 2 clk after A, 3 clk after sync1, 4 clk before C/D, 5 clk after D, 6 clk end, 7 globaltimer end
names = ["wait+A", "sync1", "issue/np", "C+D", "sync_end"]
valid = a[:, :, 7] > 0
t0 = a[:, :, 0][valid].min()
print("grid", grid, "tiles timed", valid.sum())
for it in range(4):
    v = a[:, it, :]
    ok = v[:, 7] > 0
    if not ok.any(): break
    d = np.diff(v[ok, 1:7], axis=1)
    print(f"iter {it}: ctas {ok.sum():4d} start ns p10/med/max {np.percentile(v[ok,0]-t0,10):6.0f}/{np.median(v[ok,0]-t0):6.0f}/{(v[ok,0]-t0).max():6.0f} dur ns med {np.median(v[ok,7]-v[ok,0]):6.0f} cyc " +
          " ".join(f"{n}={np.median(d[:,i]):5.0f}/{d[:,i].max():5.0f}" for i, n in enumerate(names)))
print("span of timed tiles (ns):", a[:, :, 7][valid].max() - t0)
