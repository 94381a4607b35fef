#!/usr/bin/env python
"""Benchmark of the gpushare scheduling hot path on B200 (contract: see the task statement).

  python bench.py --gpus N --steps K --warmup W            # product arm (CUDA, through the C ABI)
  python bench.py --impl reference --steps K --warmup W    # reference arm: the CPU port of the Go path

A step is one batched Filter (K3) of P pending pods against every node of the rank's cluster.
N=1 workload = BASELINE.json config 4 on one GPU: 50k nodes x 8 devices, 1M running pods, 10k pending
pods with `adversarial` requests (uniform in [1, 184320] MiB: ~P distinct rows, the hard case).

The K-step timed loop is repeated (--repeats, default 15); `ms_per_step` is the median of the repeats and the
spread is printed beside it.

N>1: the headline `value` is weak scaling (every rank filters the same 10k pending pods against its own
50k-node cluster, no collective on the data path).  The same JSON line carries `strong_c4`: config 4 as
BASELINE.json words it -- the ONE 50k-node cluster split over the N GPUs (contiguous node ranges), with the
gather of the shard results inside the timed region: kernel only, kernel + device-to-device all-gather,
kernel + gather + copy to one host, and the compact form of the same.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import shutil
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "pod x node Filter evals/sec at 50k nodes x 8 devices x 10k pending pods"
UNIT = "evals/s"
WORKLOAD = "C4"
REQ_KIND = "adversarial"
N_OUT_BUFFERS = 4  # rotating outputs: 4 x 62.7 MB > 126 MB L2, so no step finds its lines resident


def _gsx():
    m = importlib.import_module("gpushare-scheduler-extender_b200")
    sys.modules.setdefault("gsx", m)
    return m


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu --set full summaries: the steady-state
    capture (rotating outputs, cache control off) when there is one, else the isolated cold launch of round 1."""
    for name in ("r02_expand_steady_ncu_summary.json", "expand_ncu_summary.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            try:
                v = json.load(open(p)).get("dram_bytes_per_launch")
                if v is not None:
                    return v, name
            except Exception:
                pass
    return None, None


def _spread(xs):
    xs = sorted(xs)
    return {"median": statistics.median(xs), "min": xs[0], "max": xs[-1], "n": len(xs)}


def _pct(xs, q):
    xs = sorted(xs)
    if not xs:
        return None
    return xs[min(len(xs) - 1, int(round(q * (len(xs) - 1))))]


class ClockSampler:
    """Polls NVML for SM clock and throttle reasons while the timed regions run."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=1)
        med = statistics.median(self.samples) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# CPU side: the oracle port of the reference's Go path and the "fair" flat-array baseline
# (bench.py may execute oracle/ only here)
# ------------------------------------------------------------------------------------------------
def _oracle_cluster(sc):
    from gsx import synth
    from oracle import oracle as O

    phase = np.array([O.PENDING, O.RUNNING, O.SUCCEEDED, O.FAILED], dtype=np.uint8)[sc.pod_phase]
    c = O.Cluster()
    c.bulk_load(sc.node_total, sc.node_count, sc.node_listed, sc.pod_node, sc.pod_gpu, sc.pod_mem, phase)
    names = [synth.node_name(i) for i in range(sc.n_nodes)]
    arr = (O.C.c_char_p * len(names))(*[n.encode() for n in names])
    return c, names, arr


def cpu_filter_rate(sc, reqs, n_pods_sample: int, threads: int, repeats: int = 1):
    """Times Predicate.Handler for `n_pods_sample` pods x all nodes on `threads` host threads."""
    from oracle import oracle as O

    c, names, arr = _oracle_cluster(sc)
    pods = [O.Pod(f"pend-{i}", uid=f"uid-pend-{i}", limits=[int(r)]) for i, r in enumerate(reqs[:n_pods_sample])]
    c.filter_many(pods[: max(1, threads)], names, threads=threads, names_arr=arr)  # warm: lazily builds NodeInfos
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        c.filter_many(pods, names, threads=threads, names_arr=arr)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return len(pods) * sc.n_nodes / best, best


def cpu_fair_rate(sc, reqs_u64, stride: int, repeats: int = 5):
    """B-fair: the same decisions on flat arrays, AVX2 + OpenMP over all host threads (oracle/fair_baseline.c),
    whole C4 batch, best of `repeats` after one warm-up pass."""
    from oracle import fair as F
    from oracle import soa_reference as S

    recs = sc.gs_pod_records()
    ndev = sc.dev_count()
    recs["flags"][~S.legal(recs, ndev)] = 0
    used = S.used_from_records(recs, sc.n_nodes, sc.max_devs, ndev)
    cap, cls = sc.dev_capacity(), sc.node_class()
    out = np.zeros((len(reqs_u64), stride), dtype=np.uint32)
    F.filter_bitmap(cap, used, ndev, cls, reqs_u64, stride, out)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        F.filter_bitmap(cap, used, ndev, cls, reqs_u64, stride, out)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return {"value": len(reqs_u64) * sc.n_nodes / best, "unit": UNIT, "cores": F.threads(), "kind": "port-fair",
            "seconds": best, "avx2": F.has_avx2(),
            "sample": f"all {len(reqs_u64)} pending pods x all {sc.n_nodes} nodes, best of {repeats}; flat arrays, per-node key once per batch, "
                      "one compare per (pod, node), OpenMP over rows",
            "popcount": int(np.unpackbits(out.view(np.uint8)).sum())}


def go_probe():
    """the reference's own Go path is timed only where a Go toolchain exists."""
    exe = shutil.which("go")
    if not exe:
        return "go toolchain absent (go not on PATH); the Go reference cannot be built or timed on this box"
    try:
        return subprocess.run([exe, "version"], capture_output=True, text=True, timeout=10).stdout.strip()
    except Exception as ex:
        return f"go present but not runnable: {ex!r}"


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  The Go sources cannot be
    built here (no Go toolchain; vendored beego/logs missing), so this is the oracle port (kind "port"),
    with its cost structure, on every host core, each step a bounded sample of the C4 workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    _gsx()
    from gsx import synth

    cores = os.cpu_count() or 1
    sc = synth.make_config(WORKLOAD)
    reqs = synth.make_requests(synth.CONFIGS[WORKLOAD]["pending"], REQ_KIND, synth.CONFIGS[WORKLOAD]["seed"])
    from oracle import oracle as O

    c, names, arr = _oracle_cluster(sc)
    # pods per step: one Filter request (x 50k nodes) per core, a few more when the run is short
    per_step = cores * max(1, min(4, 100 // max(1, args.steps + args.warmup)))
    pods = [O.Pod(f"pend-{i}", uid=f"uid-pend-{i}", limits=[int(r)]) for i, r in enumerate(reqs[: per_step * 4])]
    times = []
    for it in range(args.warmup + args.steps):
        batch = [pods[(it * per_step + k) % len(pods)] for k in range(per_step)]
        t0 = time.perf_counter()
        c.filter_many(batch, names, threads=cores, names_arr=arr)
        dt = time.perf_counter() - t0
        if it >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = per_step * sc.n_nodes * len(times) / total
    sample = f"{per_step} Filter requests (pods) x {sc.n_nodes} nodes per step, of the {len(reqs)}-pod batch"
    words = (sc.n_nodes + 31) // 32
    stride = (words + 31) // 32 * 32
    try:
        fair = cpu_fair_rate(sc, reqs.astype(np.uint64), stride)
    except Exception as ex:  # report, do not hide
        fair = {"error": repr(ex)}
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": _config(sc, len(reqs), stride, 1, reference=True),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "cpu_baseline_fair": fair,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "go": go_probe(),
        "note": "oracle port of the Go handlers (logging compiled out), with the reference's cost structure: maps, locks, annotation strings "
                "re-parsed per evaluation, the cache-wide exclusive lock of GetNodeInfo (pkg/cache/cache.go:136) -- it does not scale with "
                "cores.  cpu_baseline_fair is what a flat-array SIMD implementation of the same decisions does on the same cores.",
    }
    print(json.dumps(line))
    return 0


def _config(sc, P, stride, world, reference=False):
    cfg = {"workload": WORKLOAD, "nodes_per_gpu": sc.n_nodes, "devices_per_node": sc.max_devs, "running_pods_per_gpu": sc.n_pods,
           "pending_pods": P, "requests": REQ_KIND}
    if reference:
        return cfg
    cfg.update({
        "bitmap_row_stride_words": stride,
        "l2": f"{N_OUT_BUFFERS} rotating output buffers of {P * stride * 4 / 1e6:.1f} MB (> 126 MB L2 in total)",
        "pipelining": "GS_OPT_PIPELINE_FILTERS: consecutive steps overlap on the GPU (prep of step k+1 beside expand of step k, "
                      "expand of step k+1 fills the SM slots step k frees); every step runs all of its kernels",
        "parallelism": f"headline: {world} GPU(s), one 50k-node cluster each (weak scaling, no data-path collective); "
                       "strong_c4: one 50k-node cluster split over the GPUs with the gather inside the timed region"})
    return cfg


# ------------------------------------------------------------------------------------------------
# product arm
# ------------------------------------------------------------------------------------------------
class Timer:
    """K-step loops timed with CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks."""

    def __init__(self, torch, dist, dev):
        self.torch, self.dist, self.dev = torch, dist, dev

    def fence(self):
        self.torch.cuda.synchronize()
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        if not self.dist:
            return v
        t = self.torch.tensor([v], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def loop_ms(self, fn, steps, repeats, warmup=0):
        """[ms per step] for `repeats` timed loops of `steps` calls of fn(i)."""
        torch = self.torch
        for i in range(warmup):
            fn(i)
        out = []
        for _ in range(repeats):
            self.fence()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                fn(i)
            e1.record()
            self.fence()
            out.append(self.max_over_ranks(e0.elapsed_time(e1)) / steps)
        return out

    def wall_ms(self, fn, steps, repeats, warmup=0):
        """The same with the host clock, for calls that end synchronised (host-buffer entry points)."""
        for i in range(warmup):
            fn(i)
        out = []
        for _ in range(repeats):
            self.fence()
            t0 = time.perf_counter()
            for i in range(steps):
                fn(i)
            dt = time.perf_counter() - t0
            self.fence()
            out.append(self.max_over_ranks(1e3 * dt) / steps)
        return out


def churn_replay(gsx, snap, sc, stride, first_slot, duration_s=2.0):
    """Config 5 as SURVEY.md 8(d) words it: every 10 ms a batch of 1000 pod events (55 % add, 45 % delete of a live
    slot), every 50 ms a Filter of 1000 pending pods against all nodes followed by 200 Binds onto nodes the Filter
    passed.  Paced by the host clock; profiling off.  Latency = duration of the C-ABI call (the Filter and the Bind
    return their results; the event batch is queued asynchronously, its `applied` latency is sampled with a
    synchronise on every tenth batch)."""
    from gsx import synth

    rng = np.random.default_rng(0xC5)
    N = sc.n_nodes
    n_batches = int(duration_s / 0.010)
    region = 100_000
    live = np.zeros(region, dtype=bool)
    batches = []
    for _ in range(n_batches):
        ev = np.zeros(1000, dtype=synth.EVENT_DT)
        slots = rng.choice(region, 1000, replace=False)
        is_add = (rng.random(1000) < 0.55) | ~live[slots]
        ev["slot"] = first_slot + slots
        ev["op"] = (~is_add).astype(np.uint8)
        ev["node_idx"] = rng.integers(0, N, 1000)
        ev["dev"] = rng.integers(0, sc.max_devs, 1000)
        ev["mem"] = rng.choice(synth.CATALOGUE[:6], 1000)
        ev["counted"] = 1
        live[slots] = is_add
        batches.append(ev)
    bind_slot = first_slot + region
    creq = gsx.PinnedBuffer(1000, np.uint64)
    creq.array[:] = synth.make_requests(1000, REQ_KIND, 0xC5).astype(np.uint64)
    cout = gsx.PinnedBuffer((1000, stride), np.uint32)
    crank = gsx.PinnedBuffer((1, N), np.uint16)
    ccut = gsx.PinnedBuffer(1000, np.uint16)
    # warm every path once (staging buffers, first-launch costs)
    snap.pods_apply_async(batches[0][:0])
    snap.filter_batch(creq.array, stride, out=cout.array)
    snap.filter_batch_compact(creq.array, out_rank=crank.array, out_cutoff=ccut.array)
    lat = {"event_batch_enqueue": [], "event_batch_applied": [], "filter_1000": [], "filter_1000_compact": [], "bind_200": []}
    lag = []
    n_ev = n_f = n_b = bound = 0
    t0 = time.perf_counter()
    for k, ev in enumerate(batches):
        due = t0 + k * 0.010
        now = time.perf_counter()
        if due - now > 0.002:
            time.sleep(due - now - 0.001)
        while time.perf_counter() < due:
            pass
        lag.append(time.perf_counter() - due)
        a = time.perf_counter()
        snap.pods_apply_async(ev)
        b = time.perf_counter()
        lat["event_batch_enqueue"].append(b - a)
        if k % 10 == 9:
            snap.sync()
            lat["event_batch_applied"].append(time.perf_counter() - a)
        n_ev += len(ev)
        if k % 5 == 4:
            a = time.perf_counter()
            snap.filter_batch(creq.array, stride, out=cout.array)
            lat["filter_1000"].append(time.perf_counter() - a)
            a = time.perf_counter()
            snap.filter_batch_compact(creq.array, out_rank=crank.array, out_cutoff=ccut.array)
            lat["filter_1000_compact"].append(time.perf_counter() - a)
            # kube-scheduler's choice is not part of the path: here, the first passing node at or after a pseudo-random
            # position of the row (binds spread over the cluster; a queue of binds on ONE node is serial by definition)
            rows = cout.array[:200]
            W = rows.shape[1]
            start = (np.arange(200, dtype=np.int64) * 2654435761 + k * 40503) % W
            cols = (np.arange(W, dtype=np.int64)[None, :] + start[:, None]) % W
            rolled = np.take_along_axis(rows, cols, axis=1)
            first_roll = (rolled != 0).argmax(axis=1)
            first_word = cols[np.arange(200), first_roll]
            w = rows[np.arange(200), first_word]
            ok = w != 0
            low = (w & (~w + np.uint32(1))).astype(np.uint32)
            bit = np.where(ok, np.log2(np.maximum(low, 1)).astype(np.int64), 0)
            nodes = (first_word * 32 + bit)[ok]
            binds = synth.bind_records(nodes, creq.array[:200][ok].view(np.int64), first_slot=bind_slot)
            bind_slot += len(binds)
            a = time.perf_counter()
            got = snap.bind_batch(binds)
            lat["bind_200"].append(time.perf_counter() - a)
            n_f += 1
            n_b += len(binds)
            bound += int((got >= 0).sum())
    snap.sync()
    wall = time.perf_counter() - t0
    # unpaced: the same event batches back to back, each waited for (what the snapshot can absorb)
    t1 = time.perf_counter()
    for ev in batches[:100]:
        snap.pods_apply(ev)
    sat = 100 * 1000 / (time.perf_counter() - t1)
    t1 = time.perf_counter()
    for ev in batches[100:200]:
        snap.pods_apply_async(ev)
    snap.sync()
    sat_async = 100 * 1000 / (time.perf_counter() - t1)
    res = {"schedule": "1000 events / 10 ms; Filter P=1000 + Bind B=200 / 50 ms", "duration_s": wall, "events": n_ev,
           "events_per_sec_sustained": n_ev / wall, "filters_of_1000": n_f, "binds": n_b, "binds_placed": bound,
           "schedule_lag_ms": {"p50": 1e3 * _pct(lag, 0.5), "p99": 1e3 * _pct(lag, 0.99), "max": 1e3 * max(lag)},
           "latency_us": {k: {"p50": 1e6 * _pct(v, 0.5), "p90": 1e6 * _pct(v, 0.9), "p99": 1e6 * _pct(v, 0.99), "max": 1e6 * max(v),
                              "max_at_call": int(np.argmax(v)), "n": len(v)} for k, v in lat.items() if v},
           "events_per_sec_saturated_sync": sat, "events_per_sec_saturated_async": sat_async,
           "note": "latencies are C-ABI call durations on the host clock, profiling off; the Filter returns the [1000][N/32] bitmap in "
                   "pinned host memory (6.3 MB over PCIe), filter_1000_compact the rank/cutoff form of the same answer"}
    for b in (creq, cout, crank, ccut):
        b.free()
    return res


def strong_c4(args, gsx, torch, dist, tm, rank, world, local_rank, dev, reqs, peak):
    """BASELINE config 4 as worded: ONE 50k-node cluster, node ranges sharded over the GPUs, gather of the shard
    results inside the timed region.  evals/s = P x N_total / time, max over ranks."""
    from gsx import sharding, synth

    cfg = synth.CONFIGS[WORKLOAD]
    sc = synth.make_cluster(cfg["n_nodes"], cfg["n_devs"], cfg["dev_cap"], cfg["running"], seed=cfg["seed"])  # the same on every rank
    sh = sharding.NodeSharding(sc.n_nodes, world)
    snap = sharding.load_shard(sh, rank, sc, device=local_rank)
    snap.set_option(gsx.capi.OPT_PIPELINE_FILTERS, 1)
    P = len(reqs)
    stride = sh.stride_words
    d_req = torch.from_numpy(reqs.view(np.int64)).to(dev)
    outs = [torch.zeros((P, stride), dtype=torch.int32, device=dev) for _ in range(N_OUT_BUFFERS)]
    gathered = [torch.zeros((world * P, stride), dtype=torch.int32, device=dev) for _ in range(2)]
    h_all = torch.zeros((world * P, stride), dtype=torch.int32).pin_memory() if rank == 0 else None
    h_mine = torch.zeros((P, stride), dtype=torch.int32).pin_memory()
    d_rank = torch.zeros(sh.per_shard, dtype=torch.int16, device=dev)
    d_cut = torch.zeros(P, dtype=torch.int16, device=dev)
    g_rank = torch.zeros(world * sh.per_shard, dtype=torch.int16, device=dev)
    h_rank = torch.zeros(world * sh.per_shard, dtype=torch.int16).pin_memory()
    h_cut = torch.zeros(P, dtype=torch.int16).pin_memory()
    stream = torch.cuda.current_stream(dev).cuda_stream or 1  # (0 would mean the snapshot's own stream)

    def kernel(i):
        snap.filter_batch_device(d_req.data_ptr(), P, outs[i % N_OUT_BUFFERS].data_ptr(), stride, stream)

    def kernel_gather(i):
        kernel(i)
        dist.all_gather_into_tensor(gathered[i % 2], outs[i % N_OUT_BUFFERS])

    def kernel_gather_d2h(i):
        kernel_gather(i)
        if rank == 0:
            h_all.copy_(gathered[i % 2], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()          # the host has the answer

    def kernel_d2h_direct(i):
        kernel(i)
        h_mine.copy_(outs[i % N_OUT_BUFFERS], non_blocking=True)  # every rank's shard over its own PCIe link
        torch.cuda.current_stream(dev).synchronize()

    def compact(i):
        snap.filter_batch_compact_device(d_req.data_ptr(), P, d_rank.data_ptr(), sh.per_shard, d_cut.data_ptr(), stream)

    def compact_gather(i):
        compact(i)
        dist.all_gather_into_tensor(g_rank.view(torch.uint8), d_rank.view(torch.uint8))

    def compact_gather_d2h(i):
        compact_gather(i)
        if rank == 0:
            h_rank.copy_(g_rank, non_blocking=True)
            h_cut.copy_(d_cut, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()

    # the gather done by the rank kernel itself: every rank's k_node_rank stores its slice into all ranks' copies of the
    # gathered array (symmetric memory, plain peer stores over NVLink), then one device-side barrier
    peer = None
    try:
        import torch.distributed._symmetric_memory as symm

        bufs, ptrs, hdl = [], [], None
        for _ in range(2):
            t = symm.empty(world * sh.per_shard, dtype=torch.int16, device=dev)
            hdl = symm.rendezvous(t, dist.group.WORLD)
            t.zero_()
            bufs.append(t)
            ptrs.append([int(x) for x in hdl.buffer_ptrs])
        torch.cuda.synchronize()
        dist.barrier()
        peer = (hdl, bufs, ptrs)
    except Exception as ex:  # report, do not hide
        peer_error = repr(ex)

    def compact_peer(i):
        hdl, bufs, ptrs = peer
        snap.filter_batch_compact_device_gather(d_req.data_ptr(), P, ptrs[i & 1], rank * sh.per_shard, sh.per_shard, d_cut.data_ptr(), stream)
        hdl.barrier(channel=0)

    def compact_peer_d2h(i):
        compact_peer(i)
        if rank == 0:
            h_rank.copy_(peer[1][i & 1], non_blocking=True)
            h_cut.copy_(d_cut, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()

    K = max(5, min(args.steps, 50))
    R = max(3, min(args.repeats, 5))
    evals = P * sc.n_nodes
    out = {"workload": "C4 split over the GPUs", "nodes_total": sc.n_nodes, "nodes_per_shard": sh.per_shard,
           "shard_bitmap_bytes": P * stride * 4, "gather": "torch.distributed all_gather_into_tensor (NCCL) on device buffers, shard-major [G][P][W]",
           "steps": K, "repeats": R}
    for name, fn in (("kernel_only", kernel), ("kernel_gather", kernel_gather), ("kernel_gather_d2h_one_host", kernel_gather_d2h),
                     ("kernel_d2h_per_rank", kernel_d2h_direct), ("compact_kernel_only", compact), ("compact_gather", compact_gather),
                     ("compact_gather_d2h_one_host", compact_gather_d2h)):
        ms = tm.loop_ms(fn, K, R, warmup=3)
        sp = _spread(ms)
        out[name] = {"ms_per_step": sp["median"], "ms_min": sp["min"], "ms_max": sp["max"], "evals_per_sec": evals / (sp["median"] * 1e-3)}
    if peer is not None:
        for name, fn in (("compact_peer_gather", compact_peer), ("compact_peer_gather_d2h_one_host", compact_peer_d2h)):
            sp = _spread(tm.loop_ms(fn, K, R, warmup=4))
            out[name] = {"ms_per_step": sp["median"], "ms_min": sp["min"], "ms_max": sp["max"], "evals_per_sec": evals / (sp["median"] * 1e-3)}
        compact_gather(0)
        compact_peer(0)
        torch.cuda.synchronize()
        out["peer_gather_equals_nccl_gather"] = bool((peer[1][0] == g_rank).all().item())
        out["peer_gather"] = "k_node_rank stores its slice into every rank's copy (torch symmetric memory, peer stores over NVLink) + one symmetric-memory barrier"
    else:
        out["peer_gather_error"] = peer_error
    out["bytes_to_host_per_step"] = {"bitmap": world * P * stride * 4, "compact": world * sh.per_shard * 2 + P * 2}
    # parity spot check inside the bench run: the gathered compact form implies the gathered bitmap (first 64 pods)
    kernel_gather(0)
    compact_gather(0)
    torch.cuda.synchronize()
    g = gathered[0].view(world, P, stride)[:, :64].cpu().numpy().view(np.uint32)
    rk = g_rank.cpu().numpy().view(np.uint16).reshape(world, sh.per_shard)
    ct = d_cut.cpu().numpy().view(np.uint16)[:64]
    bits = (rk[:, None, :] >= ct[None, :, None])
    want = np.packbits(bits, axis=2, bitorder="little").view(np.uint32)
    out["compact_equals_bitmap_on_sample"] = bool((want == g[:, :, : sh.words_per_shard]).all())
    snap.close()
    return out


def run_product(args):
    import torch

    gsx = _gsx()
    from gsx import synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_

        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    cfg = synth.CONFIGS[WORKLOAD]
    sc = synth.make_cluster(cfg["n_nodes"], cfg["n_devs"], cfg["dev_cap"], cfg["running"], seed=cfg["seed"] + 1000 * rank)
    P = cfg["pending"]
    reqs = synth.make_requests(P, REQ_KIND, cfg["seed"]).astype(np.uint64)
    N = sc.n_nodes
    words = (N + 31) // 32
    stride = (words + 31) // 32 * 32  # 128-byte rows
    R = max(1, args.repeats)

    snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + P + 140_000, device=local_rank)
    snap.node_upsert(sc.gs_nodes())
    snap.pods_load(sc.gs_pod_records())
    snap.rebuild()
    # consecutive batches may overlap on the GPU (prep of batch k+1 under the tail of expand of batch k); the
    # request array is resident before the first step and the four outputs rotate, as the option requires
    snap.set_option(gsx.capi.OPT_PIPELINE_FILTERS, 1)

    dev = torch.device("cuda", local_rank)
    d_req = torch.from_numpy(reqs.view(np.int64)).to(dev)
    outs = [torch.empty((P, stride), dtype=torch.int32, device=dev) for _ in range(N_OUT_BUFFERS)]
    # a real (non-default) stream: the library treats a NULL stream as "the snapshot's own stream", and
    # torch's events below must be recorded on the stream the kernels run on
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0
    tm = Timer(torch, dist, dev)

    def step(i):
        snap.filter_batch_device(d_req.data_ptr(), P, outs[i % N_OUT_BUFFERS].data_ptr(), stride, stream)

    sampler = ClockSampler(local_rank)
    sampler.start()

    # ---- headline: K pipelined steps, repeated; inputs resident in HBM ----
    for i in range(args.warmup):
        step(i)
    tm.fence()
    launches1 = snap.launch_count
    step_ms = tm.loop_ms(step, args.steps, R)
    gpu_launches = snap.launch_count - launches1
    sp = _spread(step_ms)
    ms_per_step = sp["median"]
    value = world * P * N / (ms_per_step * 1e-3)

    # ---- per-kernel device times: loops of ONE stage (same inputs, same rotating outputs) ----
    kernel_ms, kernel_ms_spread = {}, {}
    for name, mask in (("prep_pods", gsx.capi.STAGE_SORT), ("prep_nodes", gsx.capi.STAGE_NODE_RANK),
                       ("prep", gsx.capi.STAGE_SORT | gsx.capi.STAGE_NODE_RANK), ("expand", gsx.capi.STAGE_EXPAND)):
        def stage(i, mask=mask):
            snap.filter_stages_device(mask, d_req.data_ptr(), P, outs[i % N_OUT_BUFFERS].data_ptr(), stride, stream)
        s_ = _spread(tm.loop_ms(stage, args.steps, max(3, min(R, 9)), warmup=3))
        kernel_ms[name], kernel_ms_spread[name] = s_["median"], s_
    t_expand = kernel_ms.get("expand")
    words_pad = (words + 127) // 128 * 128
    expand_bytes = P * words * 4 + words_pad * 40 + P * 12      # bitmap written + node bucket planes + per-row request/perm/bucket
    step_bytes = N * sc.max_devs * 8 + N * 8 + N * 4 + P * 8 + P * words * 4
    peak, peak_src = _peaks()
    # k_expand is launched once per step; in the timed region consecutive launches overlap, so what a launch costs there
    # is the step period (which also carries k_prep: charging all of it to k_expand understates the kernel).  The same
    # kernel timed alone -- loops of expand-only launches, no overlap, its ramp and drain exposed -- is reported beside it.
    achieved = expand_bytes / (ms_per_step * 1e-3) / 1e9
    achieved_alone = expand_bytes / (t_expand * 1e-3) / 1e9 if t_expand else None
    traffic, traffic_src = _traffic()
    roofline = {"bound": "hbm", "kernel": "k_expand", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": expand_bytes,
                "duration_ms": ms_per_step,
                "duration_source": "timed region: CUDA events around K steps on the launching stream / K (one k_expand launch per "
                                   "step, consecutive launches overlapping; the period also contains k_prep)",
                "achieved_alone": achieved_alone, "frac_alone": (achieved_alone / peak) if achieved_alone else None,
                "kernel_ms": t_expand, "kernel_ms_spread": kernel_ms_spread.get("expand"),
                "note": "kernel_ms / frac_alone: median of repeated loops of expand-only launches into four rotating 62.7 MB outputs "
                        "(every line reaches DRAM), launches not overlapping",
                "step_algorithmic_bytes": step_bytes,
                "step_achieved_gbs": step_bytes / (ms_per_step * 1e-3) / 1e9,
                "step_frac": step_bytes / (ms_per_step * 1e-3) / 1e9 / peak}

    extra = {}
    # ---- bind (K4) and rebuild (K1) rates, device-timed through the library's per-kernel counters ----
    try:
        bits = gsx.unpack_bitmap(outs[0][:2048].cpu().numpy().view(np.uint32), N)
        B = min(2048, P)
        # bind target = a pseudo-random passing node per pod (kube-scheduler's pick is not part of the path)
        cnt = bits.sum(axis=1)
        ok = cnt > 0
        pick = (np.arange(B, dtype=np.int64) * 2654435761 % np.maximum(cnt, 1))
        csum = np.cumsum(bits, axis=1)
        first = (csum > pick[:, None]).argmax(axis=1)
        binds = synth.bind_records(first[ok], reqs[:B][ok].view(np.int64), first_slot=sc.n_pods)
        snap.profile(True)
        snap.profile_reset()
        t0 = time.perf_counter()
        got = snap.bind_batch(binds)
        t_bind_wall = time.perf_counter() - t0
        snap.bind_rollback(binds["slot"][got >= 0])
        for _ in range(9):
            snap.rebuild()
        pr = snap.profile_read()
        snap.profile(False)
        bind_ms = (pr["bind_sort"][1] + pr["bind_scan"][1])
        extra["binds_per_sec_device"] = len(binds) / (bind_ms * 1e-3) if bind_ms else None
        extra["binds_per_sec_e2e"] = len(binds) / t_bind_wall
        extra["bind_batch"] = len(binds)
        rb = pr["rebuild"]
        n_calls = rb[0] // 2  # zero + accumulate per call
        rb_bytes = (sc.n_pods + P + 140_000) * 16 + N * sc.max_devs * 8
        extra["rebuild_ms"] = rb[1] / n_calls if n_calls else None
        extra["rebuild_gbs"] = rb_bytes / (extra["rebuild_ms"] * 1e-3) / 1e9 if n_calls else None
        extra["rebuild_frac_of_hbm"] = extra["rebuild_gbs"] / peak if n_calls else None
        extra["rebuild_note"] = "per call = zeroing of used[] + k_rebuild over the whole pod table (device time between CUDA events)"
    except Exception as ex:  # report, do not hide
        extra["bind_error"] = repr(ex)

    # ---- end to end through the host-buffer C-ABI calls: pinned host buffers in and out, every step ----
    h_req = gsx.PinnedBuffer(P, np.uint64)
    h_req.array[:] = reqs
    h_out = gsx.PinnedBuffer((P, stride), np.uint32)
    h_rank = gsx.PinnedBuffer((1, N), np.uint16)
    h_cut = gsx.PinnedBuffer(P, np.uint16)
    e2e_steps = max(3, min(args.steps, 20))
    e2e_rep = max(3, min(R, 7))
    sp_e = _spread(tm.wall_ms(lambda i: snap.filter_batch(h_req.array, stride, out=h_out.array), e2e_steps, e2e_rep, warmup=2))
    e2e = {"value": world * P * N / (sp_e["median"] * 1e-3), "unit": UNIT, "h2d_bytes_per_step": P * 8, "d2h_bytes_per_step": P * stride * 4,
           "ms_per_step": sp_e["median"], "ms_min": sp_e["min"], "ms_max": sp_e["max"], "steps": e2e_steps, "repeats": e2e_rep,
           "api": "gs_filter_batch (pinned host buffers): requests in, the whole [P][N/32] bitmap out"}
    sp_c = _spread(tm.wall_ms(lambda i: snap.filter_batch_compact(h_req.array, out_rank=h_rank.array, out_cutoff=h_cut.array),
                              e2e_steps, e2e_rep, warmup=2))
    compact_ok = bool((gsx.expand_compact(h_rank.array, h_cut.array[:256], N, stride)[:256] == h_out.array[:256]).all())
    e2e_compact = {"value": world * P * N / (sp_c["median"] * 1e-3), "unit": UNIT, "h2d_bytes_per_step": P * 8,
                   "d2h_bytes_per_step": N * 2 + P * 2, "ms_per_step": sp_c["median"], "ms_min": sp_c["min"], "ms_max": sp_c["max"],
                   "steps": e2e_steps, "repeats": e2e_rep, "expands_to_the_bitmap_on_sample": compact_ok,
                   "api": "gs_filter_batch_compact: rank[N] + cutoff[P] (u16), pass(p, n) <=> rank[n] >= cutoff[p]; "
                          "tests/test_gpu_compact_subset.py proves the expansion equal to the bitmap bit for bit"}
    # one Filter request as kube-scheduler sends it: one pod, a candidate list
    sub_lat = {}
    rng = np.random.default_rng(1)
    for C in (30, 5000):
        sub = rng.choice(N, C, replace=False).astype(np.uint32)
        one = reqs[:1].copy()
        ts = []
        for it in range(60):
            a = time.perf_counter()
            snap.filter_batch_subset(one, sub)
            if it >= 10:
                ts.append(time.perf_counter() - a)
        sub_lat[f"1_pod_x_{C}_candidates_us"] = {"p50": 1e6 * _pct(ts, 0.5), "p99": 1e6 * _pct(ts, 0.99)}
    e2e_subset = {"latency": sub_lat, "api": "gs_filter_batch_subset (pkg/scheduler/predicate.go:21: the candidate list of one request)"}

    # the same compact call from three host threads at once, each with its own page-locked buffers (the extender's HTTP
    # handlers are concurrent; the library gives each caller a filter context of its own): calls per second of all of them
    def _concurrent(n_threads, steps):
        bufs = []
        for _ in range(n_threads):
            r = gsx.PinnedBuffer(P, np.uint64)
            r.array[:] = reqs
            bufs.append((r, gsx.PinnedBuffer((1, N), np.uint16), gsx.PinnedBuffer(P, np.uint16)))
        start = threading.Barrier(n_threads + 1)
        errs = []

        def work(b):
            try:
                for _ in range(3):
                    snap.filter_batch_compact(b[0].array, out_rank=b[1].array, out_cutoff=b[2].array)
                start.wait()
                for _ in range(steps):
                    snap.filter_batch_compact(b[0].array, out_rank=b[1].array, out_cutoff=b[2].array)
            except Exception as ex:  # report, do not hide
                errs.append(repr(ex))
                try:
                    start.abort()
                except Exception:
                    pass

        ths = [threading.Thread(target=work, args=(b,)) for b in bufs]
        for t in ths:
            t.start()
        start.wait()
        t0 = time.perf_counter()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        same = all(bool((b[1].array == h_rank.array).all() and (b[2].array == h_cut.array).all()) for b in bufs)
        for b in bufs:
            for x in b:
                x.free()
        return dt, same, errs

    try:
        n_thr = 3
        dts = []
        same_all = True
        for _ in range(e2e_rep):
            tm.fence()
            dt, same, errs = _concurrent(n_thr, e2e_steps * 5)
            if errs:
                raise RuntimeError("; ".join(errs))
            same_all &= same
            dts.append(tm.max_over_ranks(1e3 * dt) / (n_thr * e2e_steps * 5))
        sp_t = _spread(dts)
        extra["e2e_compact_concurrent"] = {"value": world * P * N / (sp_t["median"] * 1e-3), "unit": UNIT, "threads": n_thr,
                                           "ms_per_call_aggregate": sp_t["median"], "ms_min": sp_t["min"], "ms_max": sp_t["max"],
                                           "calls_per_thread": e2e_steps * 5, "repeats": e2e_rep, "same_answer_in_every_thread": same_all,
                                           "h2d_bytes_per_step": P * 8, "d2h_bytes_per_step": N * 2 + P * 2}
    except Exception as ex:  # report, do not hide
        extra["e2e_compact_concurrent"] = {"error": repr(ex)}

    # ---- config 5: churn replay on the paced schedule (single GPU) ----
    if world == 1 and not args.no_churn:
        try:
            extra["churn"] = churn_replay(gsx, snap, sc, stride, first_slot=sc.n_pods + P + 8192)
        except Exception as ex:
            extra["churn_error"] = repr(ex)

    gather = None
    strong = None
    if dist:
        # what a single consumer of all ranks' 62.7 MB bitmaps would pay (weak case): all-gather of [P][stride] per rank
        big = torch.empty((world * P, stride), dtype=torch.int32, device=dev)
        ms = tm.loop_ms(lambda i: dist.all_gather_into_tensor(big, outs[0]), 5, 3, warmup=2)
        gather = {"bitmap_allgather_ms": statistics.median(ms), "bytes_per_rank": P * stride * 4}
        del big
        try:
            strong = strong_c4(args, gsx, torch, dist, tm, rank, world, local_rank, dev, reqs, peak)
        except Exception as ex:
            strong = {"error": repr(ex)}

    clocks = sampler.stop()

    cpu_baseline = cpu_fair = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        n_sample = 16 * cores
        v, secs = cpu_filter_rate(sc, reqs.view(np.int64), n_sample, cores)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{n_sample} of the {P} pending pods x all {N} nodes ({secs:.1f} s on {cores} threads)"}
        try:
            cpu_fair = cpu_fair_rate(sc, reqs, stride)
        except Exception as ex:
            cpu_fair = {"error": repr(ex)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "ms_per_step_spread": sp, "repeats": R, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": _config(sc, P, stride, world),
            "roofline": roofline, "cpu_baseline": cpu_baseline, "cpu_baseline_fair": cpu_fair,
            # The end-to-end headline is the compact form of the answer (round 1's review: "report it as e2e_compact
            # beside the bitmap e2e, which stays the headline until that test is green" -- tests/test_gpu_compact_subset.py
            # is green, and this run re-checked the expansion against the bitmap on a sample); the bitmap form stays
            # beside it.  If the check of this run failed, the bitmap form is the headline.
            "e2e": dict(e2e_compact, form="compact") if compact_ok else dict(e2e, form="bitmap"),
            "e2e_bitmap": e2e, "e2e_compact": e2e_compact,
            "e2e_subset": e2e_subset, "gpu_launches": int(gpu_launches), "gpu_launches_per_step": int(gpu_launches) // max(1, args.steps * R),
            "kernel_ms": kernel_ms, "clocks": clocks, "sm_count": snap.sm_count,
        }
        line.update(extra)
        if gather:
            line["gather"] = gather
        if strong:
            line["strong_c4"] = strong
        print(json.dumps(line))
    snap.close()
    if dist:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--repeats", type=int, default=15, help="how many times the K-step timed loop is repeated (median reported)")
    ap.add_argument("--impl", default="product", choices=["product", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-churn", action="store_true")
    args = ap.parse_args()
    args.warmup = max(3, args.warmup)
    if args.impl == "reference":
        return run_reference(args)
    return run_product(args)


if __name__ == "__main__":
    sys.exit(main())
