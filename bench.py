#!/usr/bin/env python
"""Benchmark of the gpushare scheduling hot path on B200 (contract: see the task statement).

  python bench.py --gpus N --steps K --warmup W            # product arm (CUDA, through the C ABI)
  python bench.py --impl reference --steps K --warmup W    # reference arm: the CPU port of the Go path

A step is one batched Filter (K3) of P pending pods against every node of the rank's cluster shard.
N=1 workload = BASELINE.json config 4 on one GPU: 50k nodes x 8 devices, 1M running pods, 10k pending
pods with `adversarial` requests (uniform in [1, 184320] MiB: ~P distinct rows, the hard case).
N>1: weak scaling -- every rank owns its own 50k-node shard of an N x 50k-node cluster and filters the
same 10k pending pods against it; no collective on the data path (the bitmap gather a single consumer
would need is timed separately and reported under "gather").
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
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
    """dram bytes per launch of the dominant kernel from the committed ncu --set full summary, if any."""
    p = os.path.join(ROOT, "profiles", "expand_ncu_summary.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


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
# CPU side: the oracle port of the reference's Go path (bench.py may execute oracle/ only here)
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
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "nodes_per_gpu": sc.n_nodes, "devices_per_node": sc.max_devs,
                   "running_pods_per_gpu": sc.n_pods, "pending_pods": len(reqs), "requests": REQ_KIND, "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "oracle port of the Go handlers (logging compiled out); Go toolchain absent so the reference itself cannot run",
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
# product arm
# ------------------------------------------------------------------------------------------------
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

    snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + P + 8192, device=local_rank)
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

    def step(i):
        snap.filter_batch_device(d_req.data_ptr(), P, outs[i % N_OUT_BUFFERS].data_ptr(), stride, stream)

    sampler = ClockSampler(local_rank)
    sampler.start()

    launches0 = snap.launch_count
    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    launches1 = snap.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    gpu_launches = snap.launch_count - launches1
    ms_total = e0.elapsed_time(e1)
    if dist:
        t = torch.tensor([ms_total], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    value = world * P * N / (ms_per_step * 1e-3)

    # per-kernel device times: CUDA events on the launching stream around `steps` back-to-back launches
    # of ONE stage (same inputs, same rotating outputs); the stages of a step sum to ~ms_per_step
    kernel_ms = {}
    for name, mask in (("prep_pods", gsx.capi.STAGE_SORT), ("prep_nodes", gsx.capi.STAGE_NODE_RANK),
                       ("prep", gsx.capi.STAGE_SORT | gsx.capi.STAGE_NODE_RANK), ("expand", gsx.capi.STAGE_EXPAND)):
        for i in range(3):
            snap.filter_stages_device(mask, d_req.data_ptr(), P, outs[i % N_OUT_BUFFERS].data_ptr(), stride, stream)
        torch.cuda.synchronize()
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k0.record()
        for i in range(args.steps):
            snap.filter_stages_device(mask, d_req.data_ptr(), P, outs[i % N_OUT_BUFFERS].data_ptr(), stride, stream)
        k1.record()
        torch.cuda.synchronize()
        kernel_ms[name] = k0.elapsed_time(k1) / args.steps
    t_expand = kernel_ms.get("expand")
    words_pad = (words + 127) // 128 * 128
    expand_bytes = P * words * 4 + words_pad * 40 + P * 12      # bitmap written + node bucket planes + per-row request/perm/bucket
    step_bytes = N * sc.max_devs * 8 + N * 8 + N * 4 + P * 8 + P * words * 4
    peak, peak_src = _peaks()
    achieved = expand_bytes / (t_expand * 1e-3) / 1e9 if t_expand else None
    roofline = {"bound": "hbm", "kernel": "k_expand", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None, "traffic": _traffic(), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": expand_bytes, "kernel_ms": t_expand,
                "note": "the 62.7 MB bitmap fits the 126 MB L2, so `traffic` (ncu, one isolated launch) is far below the algorithmic bytes; "
                        "in this run four output buffers rotate so every step's lines reach DRAM; a plain streaming store of the same "
                        "rows takes 12.3 us on this GPU (profiles/NOTES.md)",
                "step_algorithmic_bytes": step_bytes,
                "step_achieved_gbs": step_bytes / (ms_per_step * 1e-3) / 1e9}

    # bind (K4) and rebuild (K1) rates, device-timed through the same counters
    extra = {}
    try:
        bits = gsx.unpack_bitmap(outs[(args.steps - 1) % N_OUT_BUFFERS][:2048].cpu().numpy().view(np.uint32), N)
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
        for _ in range(3):
            snap.rebuild()
        pr = snap.profile_read()
        snap.profile(False)
        bind_ms = (pr["bind_sort"][1] + pr["bind_scan"][1])
        extra["binds_per_sec_device"] = len(binds) / (bind_ms * 1e-3) if bind_ms else None
        extra["binds_per_sec_e2e"] = len(binds) / t_bind_wall
        extra["bind_batch"] = len(binds)
        rb = pr["rebuild"]
        extra["rebuild_ms"] = rb[1] / rb[0] if rb[0] else None
        extra["rebuild_gbs"] = (sc.n_pods * 16 + N * sc.max_devs * 8) / (extra["rebuild_ms"] * 1e-3) / 1e9 if rb[0] else None
    except Exception as ex:  # report, do not hide
        extra["bind_error"] = repr(ex)

    # config 5 (churn replay): batches of 1000 pod events (55 % add, 45 % delete of a live slot) interleaved with
    # a Filter of 1000 pods + 200 Binds every fifth batch, back to back through the host-buffer C-ABI calls
    try:
        rngc = np.random.default_rng(0xC5)
        free_slot0 = sc.n_pods + 4096
        n_churn_slots = min(P, 4096)
        live = np.zeros(n_churn_slots, dtype=bool)
        batches = []
        for _ in range(50):
            ev = np.zeros(1000, dtype=synth.EVENT_DT)
            is_add = rngc.random(1000) < 0.55
            slots = rngc.integers(0, n_churn_slots, 1000)
            is_add |= ~live[slots]
            ev["slot"] = free_slot0 + slots
            ev["op"] = (~is_add).astype(np.uint8)
            ev["node_idx"] = rngc.integers(0, N, 1000)
            ev["dev"] = rngc.integers(0, sc.max_devs, 1000)
            ev["mem"] = rngc.choice(synth.CATALOGUE[:6], 1000)
            ev["counted"] = 1
            live[slots] = is_add
            batches.append(ev)
        creq = synth.make_requests(1000, REQ_KIND, 0xC5).astype(np.uint64)
        cout = np.empty((1000, stride), dtype=np.uint32)
        cb_nodes = rngc.integers(0, N, 200)
        snap.profile(True)
        snap.profile_reset()
        t0 = time.perf_counter()
        n_ev = n_f = n_b = 0
        for k, ev in enumerate(batches):
            snap.pods_apply(ev)
            n_ev += len(ev)
            if k % 5 == 4:
                snap.filter_batch(creq, stride, out=cout)
                got = snap.bind_batch(synth.bind_records(cb_nodes, creq[:200].view(np.int64), first_slot=sc.n_pods + 2048))
                snap.bind_rollback(np.arange(200, dtype=np.uint32)[got >= 0] + sc.n_pods + 2048)
                n_f += 1
                n_b += 200
        wall = time.perf_counter() - t0
        pr = snap.profile_read()
        snap.profile(False)
        extra["churn"] = {"events": n_ev, "filters_of_1000": n_f, "binds": n_b, "wall_s": wall,
                          "events_per_sec_e2e": n_ev / wall,
                          "apply_device_us_per_1000_events": 1e3 * pr["apply"][1] / max(1, pr["apply"][0] // 2),
                          "note": "wall time includes the profiling synchronisation after every kernel"}
        snap.pods_apply(synth.remove_events(free_slot0 + np.arange(n_churn_slots)))
    except Exception as ex:
        extra["churn_error"] = repr(ex)

    # end to end through the host-buffer C-ABI call: pinned request in, pinned bitmap out, every step
    h_req = gsx.PinnedBuffer(P, np.uint64)
    h_req.array[:] = reqs
    h_out = gsx.PinnedBuffer((P, stride), np.uint32)
    e2e_steps = max(3, min(args.steps, 20))
    for _ in range(2):
        snap.filter_batch(h_req.array, stride, out=h_out.array)
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        snap.filter_batch(h_req.array, stride, out=h_out.array)
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    if dist:
        t = torch.tensor([e2e_s], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e = {"value": world * P * N / e2e_s, "unit": UNIT, "h2d_bytes_per_step": P * 8, "d2h_bytes_per_step": P * stride * 4,
           "ms_per_step": e2e_s * 1e3, "steps": e2e_steps, "api": "gs_filter_batch (pinned host buffers)"}

    gather = None
    if dist:
        # what a single consumer of all shards' bitmaps would pay: all-gather of [P][stride] per rank
        big = torch.empty((world, P, stride), dtype=torch.int32, device=dev)
        for _ in range(2):
            dist.all_gather_into_tensor(big, outs[0])
        torch.cuda.synchronize()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(5):
            dist.all_gather_into_tensor(big, outs[0])
        g1.record()
        torch.cuda.synchronize()
        t = torch.tensor([g0.elapsed_time(g1) / 5], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        gather = {"bitmap_allgather_ms": float(t.item()), "bytes_per_rank": P * stride * 4}
        del big

    clocks = sampler.stop()

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        n_sample = 16 * cores
        v, secs = cpu_filter_rate(sc, reqs.view(np.int64), n_sample, cores)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{n_sample} of the {P} pending pods x all {N} nodes ({secs:.1f} s on {cores} threads)"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "nodes_per_gpu": N, "devices_per_node": sc.max_devs, "running_pods_per_gpu": sc.n_pods,
                       "pending_pods": P, "requests": REQ_KIND, "bitmap_row_stride_words": stride,
                       "l2": f"{N_OUT_BUFFERS} rotating output buffers of {P * stride * 4 / 1e6:.1f} MB (> 126 MB L2 in total)",
                       "pipelining": "GS_OPT_PIPELINE_FILTERS: consecutive steps overlap on the GPU (prep of step k+1 beside expand of step k, "
                                     "expand of step k+1 fills the SM slots step k frees); every step runs all of its kernels",
                       "parallelism": f"nodes sharded over {world} GPU(s), weak scaling, no data-path collective"},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": int(gpu_launches),
            "kernel_ms": kernel_ms, "clocks": clocks,
        }
        line.update(extra)
        if gather:
            line["gather"] = gather
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
    ap.add_argument("--impl", default="product", choices=["product", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(3, args.warmup)
    if args.impl == "reference":
        return run_reference(args)
    return run_product(args)


if __name__ == "__main__":
    sys.exit(main())
