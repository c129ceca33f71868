#!/usr/bin/env python
"""bench.py -- Solo12 MPC QP solves/s (batched, 16-step horizon) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config {2,3,5}] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workloads (BASELINE.json `configs`):
  --config 2 (default, the configuration the metric is quoted on): 4096 trot instances PER GPU with randomised states /
             velocity references / footholds, weak scaling.
  --config 3: 65,536 instances per GPU mixing trot / walk / bounding / pacing / static schedules (src/Gait.cpp:38-108).
  --config 5: the full control tick -- one MPC update solve + one whole-body tick (Solo12 kinematics / RNEA, InvKin, QPWBC,
             tau_ff) per robot and step -- on 1,048,576 robots IN TOTAL, sharded over the N GPUs (strong scaling).
History of every config: tick 0 = create + cold solve (untimed), then consecutive UPDATE ticks (osqp_update_A + bounds, warm
start) with the gait rolled one row per tick and the state advanced by the predicted x_f_applied[:12, 0] (closed loop, as
scripts/test_mpc.py:78).  A "step" is one update tick of the whole batch.  The closed-loop inputs of every tick come from an
untimed pre-pass of the GPU path itself; the solver state after tick 0 is snapshotted (qrw_mpc_state_save) and restored
before each measured leg, so the warm-up, the device-resident leg (`value`) and the end-to-end leg (`e2e`) replay the same
solves.

* value     : inputs already in HBM, CUDA events around each step's launches, L2 flushed between steps (256 MiB write), max
              over ranks.
* e2e       : the public host API per step -- BatchedMPC.run(host arrays) + get_latest_result() (config 5: followed by
              BatchedWbcTick.compute(host arrays) fed with the forces of that result): H2D of the step's inputs from PINNED
              host memory, the solve(s), D2H of the result, wall clock.
* roofline  : mpc_v2_kernel is latency-bound FP64 work out of shared / tensor memory, so the bound that applies is the DFMA
              pipe: achieved = algorithmic FLOPs (workloads.FLOP_PER_ADMM_ITER x measured ADMM iterations + setup) / CUDA-event
              time; peak = the DFMA microbenchmark committed under profiles/r02_fp64_bench.txt (tools/fp64_bench.cu, 34.2
              TFLOP/s; cuBLAS DGEMM 35.8), cross-checked live by qrw_fp64_peak.  `roofline_hbm` keeps the HBM line of the
              contract (35,340 algorithmic bytes per warm-started solve against MEASURED_PEAKS.json).
* cpu_baseline / --impl reference: the CPU oracle (restated reference path, OSQP-equivalent; the reference itself cannot be
              built here -- no OSQP / Eigen) with ONE WORKER PROCESS PER HOST CORE (pinned), the whole workload of one GPU
              sharded over the workers, same seeded inputs and history.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "quadruped-reactive-walking_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

UNIT = "solves/s"
BYTES_PER_SOLVE = 35340  # in 3,556 + out 3,080 + warm-start state r/w 28,704
# dram__bytes_read.sum + dram__bytes_write.sum of one mpc_v2_kernel launch over 4096 robots, ncu --set full
# (profiles/r02x_ncu_mpc_v2.txt: 88.57 MB + 47.44 MB), per robot: no re-reads beyond the algorithmic bytes
NCU_DRAM_BYTES_PER_SOLVE = (88.568320e6 + 47.444736e6) / 4096
DFMA_PEAK_TFLOPS = 34.20  # profiles/r02_fp64_bench.txt (tools/fp64_bench.cu on this pool's B200; cuBLAS DGEMM 8192^3: 35.76)
# FP64 operations one whole-body tick executes (2 x DFMA + DADD + DMUL thread instructions of one wbc_tick_kernel launch over
# 65,536 robots, ncu --set full, profiles/r02s_ncu_wbc_tick.txt: 1.727 G DFMA, 0.692 G DADD, 1.356 G DMUL): the kernel is
# instruction-bound on the FP64 pipe (45 % active), so this is the bound its fraction is quoted against
WBC_FLOP_PER_TICK = (2 * 1.7271e9 + 0.6917e9 + 1.3557e9) / 65536
ALL_GAITS = ("trot", "walk", "bounding", "pacing", "static")
CONFIGS = {
    2: dict(metric="Solo12 MPC QP solves/sec (batched, 16-step)", unit="solves/s", batch=4096, total=None, gaits=("trot",),
            scaling="weak", full_tick=False,
            workload="4096 batched Solo12 trot MPC instances per GPU, randomised states/velocity refs, update ticks (warm start)"),
    3: dict(metric="Solo12 MPC QP solves/sec (batched, 16-step)", unit="solves/s", batch=65536, total=None, gaits=ALL_GAITS,
            scaling="weak", full_tick=False,
            workload="65536 batched MPC instances per GPU mixing trot/walk/bounding/pacing/static contact schedules, update ticks"),
    5: dict(metric="Solo12 full control ticks/sec (MPC update solve + whole-body tick per robot)", unit="ticks/s", batch=None,
            total=1 << 20, gaits=ALL_GAITS, scaling="strong", full_tick=True,
            workload="full control tick (MPC + QPWBC + InvKin/RNEA) of 1,048,576 robots in total, sharded over the GPUs"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=0, help="instances per GPU (0 = the config's own size)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-workers", type=int, default=0, help="worker processes of the CPU arm (0 = one per host core)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="instances of the CPU arm (0 = the whole per-GPU workload, capped at 4096)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--work-order", type=int, default=1, choices=[0, 1],
                    help="1 (default, the library's default): the persistent MPC grid takes the longest-expected solves first; 0: index order (A/B)")
    ap.add_argument("--wbc-batch", type=int, default=-1,
                    help="robots in the auxiliary whole-body-tick / planner legs (BASELINE config 4); 0 skips them; default 65536 for config 2")
    return ap.parse_args()


def per_gpu_batch(args, world):
    cfg = CONFIGS[args.config]
    if args.batch > 0:
        return args.batch
    return cfg["batch"] if cfg["batch"] else cfg["total"] // world


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle port, one pinned worker process per host core (spawned: no CUDA state is inherited)
# ------------------------------------------------------------------------------------------
def _cpu_worker(idx, workers, cores, B, seed, gaits, ticks, bar):
    try:
        os.sched_setaffinity(0, {cores[idx % len(cores)]})
    except Exception:
        pass
    from oracle import bindings as ob
    from qrw_b200 import fleet, workloads as wl

    lo, hi = fleet.shard_bounds(B, idx, workers)
    w = wl.MpcWorkload(B, seed, gaits=gaits, ticks=ticks + 2)
    for name in ("x0", "vref", "footholds", "timeline"):  # this worker's robots of the SAME seeded workload
        setattr(w, name, getattr(w, name)[lo:hi])
    w.B = hi - lo
    o = ob.OracleMPC(w.B) if w.B else None
    x0 = None
    for t in range(ticks + 1):
        bar.wait()  # start of tick t (tick 0 = cold start)
        if o:
            xr, fs = w.inputs(t, x0)
            o.run(t, xr, fs, 1)
            x0 = o.get_latest_result()[:, :12].copy()
        bar.wait()  # end of tick t


def cpu_leg(steps, warmup, B, gaits, seed=1, workers=0):
    """Times `steps` update ticks of B instances sharded over one pinned worker process per host core.
    Returns (solves_per_s, ms_per_step, cores, description)."""
    import multiprocessing as mp

    from oracle import bindings as ob

    ob.build()
    cores = sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else list(range(os.cpu_count() or 1))
    P = workers if workers > 0 else len(cores)
    P = max(1, min(P, B))
    ctx = mp.get_context("spawn")
    ticks = warmup + steps
    bar = ctx.Barrier(P + 1)
    procs = [ctx.Process(target=_cpu_worker, args=(i, P, cores, B, seed, tuple(gaits), ticks, bar), daemon=True) for i in range(P)]
    for p_ in procs:
        p_.start()
    total = 0.0
    try:
        for t in range(ticks + 1):
            bar.wait(timeout=600)
            t0 = time.perf_counter()
            bar.wait(timeout=1800)
            if t > warmup:
                total += time.perf_counter() - t0
    finally:
        for p_ in procs:
            p_.join(timeout=30)
            if p_.is_alive():
                p_.terminate()
    v = B * steps / total
    desc = (f"{B} instances of the per-GPU workload generator (same seed, gait mix and closed-loop history), tick 0 (cold) untimed, "
            f"{steps} timed update ticks after {warmup} warm-up ticks, {P} pinned worker processes ({v / P:.1f} solves/s per core)")
    return v, 1e3 * total / steps, P, desc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    B = per_gpu_batch(args, max(1, args.gpus))
    if args.cpu_sample > 0:
        B = min(B, args.cpu_sample)
    elif args.config != 2:
        B = min(B, 4096)  # bounded sample of the larger configs (same seed: the first 4096 robots)
    v, ms, cores, desc = cpu_leg(args.steps, args.warmup, B, cfg["gaits"], workers=args.cpu_workers)
    note = "" if not cfg["full_tick"] else " (MPC update solve only: the whole-body tick is < 1 % of a CPU tick)"
    line = {
        "impl": "reference", "metric": cfg["metric"], "value": v, "unit": cfg["unit"], "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": cfg["scaling"], "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": cfg["workload"], "batch_per_gpu": B, "n_steps": 16, "N_gait": 20,
                   "history": "tick 0 cold (untimed) then update ticks, closed loop", "sample": desc + note},
        "cpu_baseline": {"value": v, "unit": cfg["unit"], "cores": cores, "kind": "port", "sample": desc + note,
                         "per_core": v / cores},
        "e2e": {"value": v, "unit": cfg["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------
# clocks sampled DURING the timed region (pynvml; same fields as the nvidia-smi recipe line)
# ------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop, self._t = [], set(), None, threading.Event(), None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()

    def stop(self):
        if self._t:
            self._stop.set()
            self._t.join()
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    from qrw_b200 import BatchedMPC, BatchedWbcTick, workloads as wl
    from qrw_b200._lib import load_library

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- qrw_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = load_library()
    cfg = CONFIGS[args.config]
    B, K, W = per_gpu_batch(args, world), args.steps, args.warmup
    T = W + K
    full = cfg["full_tick"]
    wbc_batch = args.wbc_batch if args.wbc_batch >= 0 else (65536 if args.config == 2 else 0)

    def pinned_like(a):
        t = torch.empty(a.shape, dtype=torch.float64).pin_memory()
        t.numpy()[...] = a
        return t

    # ---- pre-pass (untimed): closed-loop inputs of ticks 1..T, state snapshot after tick 0
    w = wl.MpcWorkload(B, 1 + rank, gaits=cfg["gaits"], ticks=T + 2)
    mpc = BatchedMPC(B, device=local)
    mpc.set_work_order(args.work_order)
    _, _, pin_r = mpc.pinned()
    xr, fs = w.inputs(0)
    mpc.run(0, xr, fs)
    snap = torch.empty(mpc.state_bytes(), dtype=torch.uint8, device=dev)
    mpc.state_save(snap)
    torch.cuda.synchronize()
    x0 = mpc.get_latest_result()[:, :12].copy()
    cold_iters = mpc.info()["iters"].copy()
    host_in, dev_in, iters = [], [], []
    lazy_host = B > (1 << 17)  # very large shards: ONE pinned pair, refilled (untimed) from the device copy before each e2e step
    for t in range(1, T + 1):
        xr, fs = w.inputs(t, x0)
        if not lazy_host or t == 1:
            host_in.append((pinned_like(xr), pinned_like(fs)))  # pinned host copies: what the end-to-end leg feeds
        dev_in.append((torch.from_numpy(xr).to(dev), torch.from_numpy(fs).to(dev)))
        mpc.run(t, dev_in[-1][0], dev_in[-1][1])
        torch.cuda.synchronize()
        x0 = mpc.get_latest_result()[:, :12].copy()
        iters.append(mpc.info()["iters"].copy())
    del w
    status = mpc.info()["status"]
    timed_iters = np.concatenate(iters[W:])
    res = mpc.result_tensor()

    tick = targs_h = targs_d = None
    if full:  # the whole-body tick of the same robots: configuration / velocity / foot goals of BASELINE config 4
        tick = BatchedWbcTick(B, device=local)
        targs_h = [pinned_like(np.ascontiguousarray(a)) for a in wl.wbc_tick_inputs(B, seed=5 + rank)]
        targs_d = [a.to(dev) for a in targs_h]

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    # two sets of gather buffers: the all_gather of step s runs on NCCL's stream while step s + 1 is being solved
    gather = [[torch.empty(B, 12, dtype=torch.float64, device=dev) for _ in range(world)] for _ in range(2)] if world > 1 else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device(t):
        mpc.run(t + 1, *dev_in[t])
        if full:  # f_cmd = x_f_mpc[12:24, 0] read in place from the MPC result (scripts/Controller.py:258,300)
            tick.compute_from_mpc(targs_d[0], targs_d[1], res, targs_d[3], targs_d[4], targs_d[5], targs_d[6])
        if world > 1:  # the one exchange of the path: first-step contact forces of every robot
            dist.all_gather(gather[0], res[:, 12:24].contiguous())  # = fleet.gather_rows for equal shards

    # ---- leg 1: device-resident inputs, CUDA events on the launching (torch current) stream
    mpc.state_load(snap)
    for t in range(W):
        step_device(t)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    launches0 = L.qrw_launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    ev_end = torch.cuda.Event(enable_timing=True)
    pending = [None, None]
    t_wall0 = time.perf_counter()
    for s in range(K):
        flush.fill_(s & 0xFF)  # L2 flush between timed iterations (256 MiB > 126 MB L2)
        ev[s][0].record()
        kev[s][0].record()
        mpc.run(W + s + 1, *dev_in[W + s])
        kev[s][1].record()  # the MPC kernel alone (the roofline's launch duration)
        if full:
            tick.compute_from_mpc(targs_d[0], targs_d[1], res, targs_d[3], targs_d[4], targs_d[5], targs_d[6])
        if world > 1:
            # The result gather is pipelined: the forces are copied out of the result buffer (the next solve overwrites it), the
            # all_gather is queued on NCCL's own stream, and this stream waits for the gather of the PREVIOUS step only -- so a
            # rank starts its next solve without waiting for the slowest rank of this one (ranks solve different robots).
            snd = res[:, 12:24].contiguous()
            if pending[s & 1] is not None:
                pending[s & 1].wait()  # the buffers of gather[s & 1] are free again (the gather of step s - 2)
            pending[s & 1] = dist.all_gather(gather[s & 1], snd, async_op=True)
            if pending[(s + 1) & 1] is not None:
                pending[(s + 1) & 1].wait()  # the gather of step s - 1 has landed
                pending[(s + 1) & 1] = None
        ev[s][1].record()
    for p_ in pending:  # the last gather belongs to the timed region
        if p_ is not None:
            p_.wait()
    ev_end.record()
    barrier()
    tail_ms = ev[K - 1][1].elapsed_time(ev_end)
    wall_dev = time.perf_counter() - t_wall0
    launches = L.qrw_launch_count() - launches0
    clocks = sampler.stop()
    step_ms = np.array([a.elapsed_time(b) for a, b in ev])
    kern_step_ms = np.array([a.elapsed_time(b) for a, b in kev])
    tot = torch.tensor([step_ms.sum() + tail_ms], dtype=torch.float64, device=dev)  # + the wait for the last gather
    rank_ms = [float(kern_step_ms.mean())]
    if world > 1:
        allk = [torch.zeros(1, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(allk, torch.tensor([kern_step_ms.mean()], dtype=torch.float64, device=dev))
        rank_ms = [float(x.item()) for x in allk]
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    total_ms = float(tot.item())
    value = world * B * K / (total_ms * 1e-3)

    # ---- leg 2: end to end through the public host API (pinned H2D, solve, D2H), wall clock
    mpc.state_load(snap)
    torch.cuda.synchronize()

    f_pin = tick_out = None
    if full:  # pinned host buffers for the tick's force input and its five outputs
        f_pin = pinned_like(np.zeros((B, 12))).numpy()
        tick_out_t = [pinned_like(np.zeros((B, 12))) for _ in range(5)]
        tick_out = [t_.numpy() for t_ in tick_out_t]

    def step_host(t):
        hx, hf = host_in[0 if lazy_host else t]
        if lazy_host:
            hx.copy_(dev_in[t][0]), hf.copy_(dev_in[t][1])
            torch.cuda.synchronize()
        t_ = time.perf_counter()
        mpc.run(t + 1, hx.numpy(), hf.numpy())
        mpc.get_latest_result(out=pin_r)
        if full:
            a = [x.numpy() for x in targs_h]
            np.copyto(f_pin, pin_r[:, 12:24])  # x_f_mpc[12:24, 0] (Controller.py:258) on the host, as the reference does
            a[2] = f_pin
            tick.compute(*a, out=tick_out)
        return time.perf_counter() - t_

    for t in range(W):
        step_host(t)
    barrier()
    e2e_s = sum(step_host(W + s) for s in range(K))
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = world * B * K / float(e2e_t.item())
    h2d = int(host_in[0][0].numel() + host_in[0][1].numel()) * 8
    d2h = int(pin_r.nbytes)
    if full:
        h2d += sum(int(x.numel()) for x in targs_h) * 8
        d2h += 5 * B * 12 * 8

    # ---- leg 3 (auxiliary, config 2 only): BASELINE config 4 (whole-body tick alone on 65536 robots), the planners, and the
    #      closed control loop at the reference cadence through qrw_tick_*
    aux = {}
    if wbc_batch > 0:
        aux = aux_legs(args, L, dev, local, rank, world, B, K, W, wbc_batch, mpc, snap, dev_in, res, flush, barrier)

    # ---- roofline of the dominant kernel
    peaks, which = measured_peaks()
    kern_ms = float(kern_step_ms.mean())
    flop = float(timed_iters.mean()) * wl.FLOP_PER_ADMM_ITER + wl.FLOP_PER_SETUP
    ach_tf = flop * B / (kern_ms * 1e-3) / 1e12
    roof = {"bound": "fp64_dfma", "achieved": ach_tf, "peak": DFMA_PEAK_TFLOPS, "unit": "TFLOP/s", "frac": ach_tf / DFMA_PEAK_TFLOPS,
            "traffic": NCU_DRAM_BYTES_PER_SOLVE * B, "kernel": "mpc_v2_kernel", "flop_per_solve": flop,
            "kernel_ms_per_launch": kern_ms,
            "peak_source": "DFMA microbenchmark tools/fp64_bench.cu on this pool's B200 (profiles/r02_fp64_bench.txt: DFMA 34.20, DMMA 37.19, "
                           "cuBLAS DGEMM 35.76 TFLOP/s); MEASURED_PEAKS.json holds no FP64 figure",
            "flop_model": "workloads.FLOP_PER_ADMM_ITER (44,408: the block-tridiagonal ADMM iteration of DESIGN.md) x measured mean "
                          "ADMM iterations + FLOP_PER_SETUP",
            "traffic_source": "dram bytes of one launch, ncu --set full (profiles/r02x_ncu_mpc_v2.txt), scaled by batch"}
    if hasattr(L, "qrw_fp64_peak"):
        import ctypes as C

        pk = C.c_double()
        if L.qrw_fp64_peak(local, C.byref(pk)) == 0:
            roof["peak_live"] = pk.value
    achieved = BYTES_PER_SOLVE * B / (kern_ms * 1e-3) / 1e9
    roof_hbm = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                "traffic": NCU_DRAM_BYTES_PER_SOLVE * B, "peak_source": which, "algorithmic_bytes_per_solve": BYTES_PER_SOLVE,
                "note": "not the applicable bound: a solve reads 35 KB and then iterates ~580 times on chip"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    cpu = None
    if not args.no_cpu_baseline:
        Bc = min(B, args.cpu_sample) if args.cpu_sample > 0 else min(B, 4096)
        v, ms, cores, desc = cpu_leg(3, 1, Bc, cfg["gaits"], workers=args.cpu_workers)
        cpu = {"value": v, "unit": cfg["unit"], "cores": cores, "kind": "port", "sample": desc, "per_core": v / cores}
    line = {
        "metric": cfg["metric"], "value": value, "unit": cfg["unit"], "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": cfg["scaling"], "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": cfg["workload"], "bench_config": args.config, "batch_per_gpu": B, "total_robots": world * B, "n_steps": 16,
                   "N_gait": 20, "history": "tick 0 cold (untimed) then update ticks, closed loop",
                   "l2": "256 MiB flush write between timed steps", "sharding": f"instances, {world} rank(s), all_gather of f0 only, pipelined with the next step's solve"},
        "e2e": {"value": e2e_value, "unit": cfg["unit"], "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "transfer": "MPC inputs and results cross PCIe inside the timed region, zero-copy: the kernel reads each robot's inputs "
                            "from the pinned host buffers and writes its result to the pinned result buffer, under the solves"},
        "gpu_launches": int(launches),
        "roofline": roof, "roofline_hbm": roof_hbm, "cpu_baseline": cpu, "clocks": clocks,
        "latency_ms": {"p50": float(np.percentile(step_ms, 50)), "p99": float(np.percentile(step_ms, 99)), "max": float(step_ms.max())},
        "admm_iters": {"mean": float(timed_iters.mean()), "p50": float(np.percentile(timed_iters, 50)),
                       "p99": float(np.percentile(timed_iters, 99)), "max": int(timed_iters.max()),
                       "cold_tick_mean": float(cold_iters.mean())},
        "mpc_kernel_ms_per_rank": rank_ms,
        "solved_frac": float((status == 1).mean()), "wall_s_device_leg": wall_dev,
    }
    line.update(aux)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def aux_legs(args, L, dev, local, rank, world, B, K, W, Bw, mpc, snap, dev_in, res, flush, barrier):
    import torch
    import torch.distributed as dist

    from qrw_b200 import BatchedWbcTick, workloads as wl

    aux = {}
    hbm = measured_peaks()[0]["hbm_gbs"]

    def max_over_ranks(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    tick = BatchedWbcTick(Bw, device=local)
    targs = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in wl.wbc_tick_inputs(Bw, seed=3 + rank)]
    for _ in range(3):
        tick.compute(*targs)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    flush.fill_(1)
    e0.record()
    for _ in range(reps):
        tick.compute(*targs)
    e1.record()
    barrier()
    wbc_ms = max_over_ranks(e0.elapsed_time(e1)) / reps
    wbc_bytes = 1240  # 712 B in + 528 B out per tick through the batched-native API
    aux["wbc_tick"] = {"value": world * Bw / (wbc_ms * 1e-3), "unit": "ticks/s", "robots_per_gpu": Bw, "ms_per_step": wbc_ms,
                       "kernels_per_step": int(getattr(tick, "kernels_per_tick", 3)),
                       "hbm_GBps_algorithmic": wbc_bytes * Bw / (wbc_ms * 1e-3) / 1e9,
                       "hbm_frac_of_measured": wbc_bytes * Bw / (wbc_ms * 1e-3) / 1e9 / hbm,
                       "roofline": {"bound": "fp64_dfma", "achieved": WBC_FLOP_PER_TICK * Bw / (wbc_ms * 1e-3) / 1e12, "peak": DFMA_PEAK_TFLOPS,
                                    "unit": "TFLOP/s", "frac": WBC_FLOP_PER_TICK * Bw / (wbc_ms * 1e-3) / 1e12 / DFMA_PEAK_TFLOPS,
                                    "kernel": "wbc_tick_kernel", "flop_per_tick": WBC_FLOP_PER_TICK,
                                    "flop_source": "executed FP64 thread instructions of one launch (ncu), see bench.py WBC_FLOP_PER_TICK"},
                       "workload": "BASELINE config 4: Solo12 kinematics/RNEA + InvKin + QPWBC + tau_ff, inputs resident in HBM"}
    del tick
    # full control tick on the MPC batch: MPC update solve, then the whole-body tick with f_cmd read in place from the result
    tickm = BatchedWbcTick(B, device=local)
    margs = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in wl.wbc_tick_inputs(B, seed=5 + rank)]

    def full_step(t):
        mpc.run(t + 1, *dev_in[t])
        tickm.compute_from_mpc(margs[0], margs[1], res, margs[3], margs[4], margs[5], margs[6])

    mpc.state_load(snap)
    for t in range(W):
        full_step(t)
    barrier()
    e0.record()
    for s_ in range(K):
        full_step(W + s_)
    e1.record()
    barrier()
    ft = max_over_ranks(e0.elapsed_time(e1))
    aux["full_tick"] = {"value": world * B * K / (ft * 1e-3), "unit": "ticks/s", "robots_per_gpu": B,
                        "workload": "BASELINE config 5 shape on the MPC batch: one MPC update solve + one whole-body tick per robot and step "
                                    "(the 1,048,576-robot run is --config 5)"}
    del tickm

    # planners alone on the config-4 batch: one kernel per control tick, 10 ticks = one MPC period
    from qrw_b200.controller import BatchedController
    from qrw_b200.planner import BatchedPlanner

    pw = wl.PlannerWorkload(Bw, 7 + rank, change_every=1000)
    pin = [[torch.from_numpy(a).to(dev) for a in pw.tick(k)] for k in range(20)]
    pl = BatchedPlanner(Bw, device=local)
    for k in range(10):
        pl.run(k, *pin[k])
    barrier()
    flush.fill_(2)
    e0.record()
    for k in range(10, 20):
        pl.run(k, *pin[k])
    e1.record()
    barrier()
    plan_ms = max_over_ranks(e0.elapsed_time(e1)) / 10
    plan_bytes = 6532  # 1,520 B read + 5,012 B written per robot and tick (xref 1,632 + fsteps 1,920 dominate)
    aux["planner_tick"] = {"value": world * Bw / (plan_ms * 1e-3), "unit": "ticks/s", "robots_per_gpu": Bw, "ms_per_step": plan_ms,
                           "kernels_per_step": 1, "hbm_GBps_algorithmic": plan_bytes * Bw / (plan_ms * 1e-3) / 1e9,
                           "hbm_frac_of_measured": plan_bytes * Bw / (plan_ms * 1e-3) / 1e9 / hbm,
                           "roofline": {"bound": "hbm", "achieved": plan_bytes * Bw / (plan_ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                                        "frac": plan_bytes * Bw / (plan_ms * 1e-3) / 1e9 / hbm, "kernel": "planner_tick_kernel",
                                        "bytes_per_tick": plan_bytes},
                           "workload": "Gait + FootstepPlanner + FootTrajectoryGenerator + StatePlanner + feet commands, inputs resident in HBM"}
    del pl, pin
    # closed control loop at the reference cadence (Controller.py:246: one MPC solve per k_mpc = 10 ticks) through the C ABI
    # qrw_tick_*: planners -> MPC -> whole-body tick, one MPC period per CUDA-graph launch; `value` with the period's inputs
    # staged on the device, `e2e` with host buffers in and out (qrw_tick_period_host: 156 B in, 288 B out per robot and tick)
    cw = wl.PlannerWorkload(B, 9 + rank, change_every=1000, yaw0=0.0, speed=0.5)
    cin = [cw.tick(k) for k in range(10 * (2 + 2 * K))]
    ctl = BatchedController(B, device=local)
    for k in range(10):
        ctl.tick(k, *cin[k])
    stage = ctl.capture_period()

    def stack(p_):
        return [np.ascontiguousarray(np.stack([cin[10 * p_ + i][j] for i in range(10)])) for j in range(4)]

    dstk = {p_: [torch.from_numpy(a).to(dev) for a in stack(p_)] for p_ in range(1, 2 + K)}

    def period(p_):
        for dst, src in zip(stage, dstk[p_]):
            dst.copy_(src)
        ctl.replay_period()

    period(1)
    barrier()
    pev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for p_ in range(2, 2 + K):
        pev[p_ - 2][0].record()
        period(p_)
        pev[p_ - 2][1].record()
    barrier()
    period_ms = np.array([a.elapsed_time(b) for a, b in pev])
    ct = max_over_ranks(float(period_ms.sum()))
    def pin(a):  # pinned host memory (the e2e contract): a torch pinned tensor behind a numpy view of the same dtype
        t = torch.empty(a.shape, dtype=torch.from_numpy(a[:0].copy()).dtype).pin_memory()
        v = t.numpy()
        v[...] = a
        pin.keep.append(t)
        return v

    pin.keep = []
    hstk = [[pin(a) for a in stack(p_)] for p_ in range(2 + K, 2 + 2 * K)]
    outs = [pin(np.empty((10, B, 12))) for _ in range(3)]
    barrier()
    t0 = time.perf_counter()
    for h_ in hstk:
        ctl.period_host(*h_, out=outs)
    ce = max_over_ranks(1e3 * (time.perf_counter() - t0))
    aux["control_loop"] = {"value": world * B * 10 * K / (ct * 1e-3), "unit": "control ticks/s", "robots_per_gpu": B,
                           "e2e": {"value": world * B * 10 * K / (ce * 1e-3), "unit": "control ticks/s",
                                   "h2d_bytes_per_period": int(sum(a.nbytes for a in hstk[0])), "d2h_bytes_per_period": int(sum(o.nbytes for o in outs))},
                           "ms_per_mpc_period": ct / K, "period_ms": [round(float(x), 3) for x in period_ms],
                           "mpc_iters_mean_last_period": float(ctl.mpc.info()["iters"].mean()),
                           "kernels_per_period": 10 * 5 + 1, "periods": K,
                           "workload": "BASELINE config 5 at the reference cadence: 10 control ticks (planners + whole-body tick) and one "
                                       "MPC update solve per robot and period, everything resident in HBM, one CUDA graph per period "
                                       "(C ABI qrw_tick_*)"}
    ctl.close()
    return aux


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
