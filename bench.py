#!/usr/bin/env python
"""bench.py -- Solo12 MPC QP solves/s (batched, 16-step horizon) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): B = 4096 trot instances per GPU with
randomised states / velocity references / footholds.  History: tick 0 = create + cold solve
(untimed), then consecutive UPDATE ticks (osqp_update_A + bounds, warm start) with the gait rolled
one row per tick and the state advanced by the predicted x_f_applied[:12, 0] (closed loop, as
scripts/test_mpc.py:78).  A "step" is one update tick of the whole batch.  The closed-loop inputs
of every tick are produced by an untimed pre-pass of the GPU path itself; the solver state after
tick 0 is snapshotted (qrw_mpc_state_save) and restored before each measured leg, so the warm-up,
the device-resident leg (`value`) and the end-to-end leg (`e2e`) replay exactly the same solves.

* value     : inputs already in HBM, CUDA events around each step's launch, L2 flushed between
              steps (256 MiB write), max over ranks.
* e2e       : BatchedMPC.run(host arrays) + get_latest_result() per step: H2D of that step's xref /
              fsteps from pinned host memory, solve, D2H of the 24x16 result, wall clock.
* roofline  : the single kernel of the step (mpc_solve_kernel) against the measured HBM peak with
              the algorithmic bytes of SURVEY.md 8d (35,340 B / warm-started solve).  The kernel is
              NOT HBM bound (an ADMM iteration streams a 38 KB factor out of shared memory ~700
              times per solve); `fp64` adds the achieved FP64 FLOP/s against a DFMA peak measured
              by the same library (qrw_fp64_peak), see DESIGN.md.
* cpu_baseline / --impl reference: the CPU oracle (restated reference path, OSQP-equivalent; the
              reference itself cannot be built here -- no OSQP/Eigen) on all host threads, same
              workload, bounded sample.
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

METRIC = "Solo12 MPC QP solves/sec (batched, 16-step)"
UNIT = "solves/s"
BYTES_PER_SOLVE = 35340  # in 3,556 + out 3,080 + warm-start state r/w 28,704
# dram__bytes_read.sum + dram__bytes_write.sum of one mpc_v2_kernel launch over 4096 robots, ncu --set full
# (profiles/r01f_ncu_mpc_v2.txt: 89.76 MB + 49.09 MB), per robot: no re-reads beyond the algorithmic bytes
NCU_DRAM_BYTES_PER_SOLVE = (89.755392e6 + 49.092096e6) / 4096
WORKLOAD = "4096 batched Solo12 trot MPC instances per GPU, randomised states/velocity refs, update ticks (warm start)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=4096, help="instances per GPU")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="instances in the CPU sample (0 = 48 per host thread)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--work-order", type=int, default=0, choices=[0, 1],
                    help="0 (default): the persistent MPC grid takes robots in index order; 1: longest-expected-solve first (A/B)")
    ap.add_argument("--wbc-batch", type=int, default=65536, help="robots in the whole-body-tick leg (BASELINE config 4); 0 skips it")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores
# ------------------------------------------------------------------------------------------
def cpu_leg(steps, warmup, sample, seed=1):
    """Times `steps` update ticks of `sample` instances on all host threads.  Returns
    (solves_per_s, ms_per_step, cores, description)."""
    from oracle import bindings as ob
    from qrw_b200 import workloads as wl

    cores = os.cpu_count() or 1
    if sample <= 0:
        sample = 48 * cores
    w = wl.MpcWorkload(sample, seed, gaits=("trot",), ticks=warmup + steps + 2)
    o = ob.OracleMPC(sample)
    xr, fs = w.inputs(0)
    o.run(0, xr, fs, cores)
    x0 = o.get_latest_result()[:, :12].copy()
    total = 0.0
    for t in range(1, warmup + steps + 1):
        xr, fs = w.inputs(t, x0)
        t0 = time.perf_counter()
        o.run(t, xr, fs, cores)
        dt = time.perf_counter() - t0
        if t > warmup:
            total += dt
        x0 = o.get_latest_result()[:, :12].copy()
    desc = (f"first {sample} instances of the same seeded workload, tick 0 (cold) untimed, "
            f"{steps} timed update ticks after {warmup} warm-up ticks, {cores} threads")
    return sample * steps / total, 1e3 * total / steps, cores, desc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    v, ms, cores, desc = cpu_leg(args.steps, args.warmup, args.cpu_sample)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": {"workload": WORKLOAD, "sample": desc},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
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

    from qrw_b200 import BatchedMPC, workloads as wl
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
    B, K, W = args.batch, args.steps, args.warmup
    T = W + K

    # ---- pre-pass (untimed): closed-loop inputs of ticks 1..T, state snapshot after tick 0
    w = wl.MpcWorkload(B, 1 + rank, gaits=("trot",), ticks=T + 2)
    mpc = BatchedMPC(B, device=local)
    mpc.set_work_order(args.work_order)
    pin_x, pin_f, pin_r = mpc.pinned()
    xr, fs = w.inputs(0)
    mpc.run(0, xr, fs)
    snap = torch.empty(mpc.state_bytes(), dtype=torch.uint8, device=dev)
    mpc.state_save(snap)
    torch.cuda.synchronize()
    x0 = mpc.get_latest_result()[:, :12].copy()
    host_in, dev_in, iters = [], [], []
    for t in range(1, T + 1):
        xr, fs = w.inputs(t, x0)
        host_in.append((xr, fs))
        dev_in.append((torch.from_numpy(xr).to(dev), torch.from_numpy(fs).to(dev)))
        mpc.run(t, dev_in[-1][0], dev_in[-1][1])
        torch.cuda.synchronize()
        x0 = mpc.get_latest_result()[:, :12].copy()
        iters.append(mpc.info()["iters"].copy())
    status = mpc.info()["status"]
    timed_iters = np.concatenate(iters[W:])

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    gather = [torch.empty(B, 12, dtype=torch.float64, device=dev) for _ in range(world)] if world > 1 else None
    res = mpc.result_tensor()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- leg 1: device-resident inputs, CUDA events on the launching (torch current) stream
    mpc.state_load(snap)
    for t in range(W):
        mpc.run(t + 1, *dev_in[t])
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    launches0 = L.qrw_launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    t_wall0 = time.perf_counter()
    for s in range(K):
        flush.fill_(s & 0xFF)  # L2 flush between timed iterations (256 MiB > 126 MB L2)
        ev[s][0].record()
        mpc.run(W + s + 1, *dev_in[W + s])
        if world > 1:  # the one exchange of the path: first-step contact forces of every robot
            dist.all_gather(gather, res[:, 12:24].contiguous())  # = fleet.gather_rows for equal shards
        ev[s][1].record()
    barrier()
    wall_dev = time.perf_counter() - t_wall0
    launches = L.qrw_launch_count() - launches0
    clocks = sampler.stop()
    step_ms = np.array([a.elapsed_time(b) for a, b in ev])
    tot = torch.tensor([step_ms.sum()], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    total_ms = float(tot.item())
    value = world * B * K / (total_ms * 1e-3)

    # ---- leg 2: end to end through the public host API (pinned H2D, solve, D2H), wall clock
    mpc.state_load(snap)
    torch.cuda.synchronize()
    for t in range(W):
        pin_x[:], pin_f[:] = host_in[t]
        mpc.run(t + 1, pin_x, pin_f)
        mpc.get_latest_result(out=pin_r)
    barrier()
    t0 = time.perf_counter()
    for s in range(K):
        pin_x[:], pin_f[:] = host_in[W + s]  # this step's inputs arrive in pinned host memory
        mpc.run(W + s + 1, pin_x, pin_f)
        mpc.get_latest_result(out=pin_r)
    e2e_s = time.perf_counter() - t0
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = world * B * K / float(e2e_t.item())
    h2d = int(pin_x.nbytes + pin_f.nbytes)
    d2h = int(pin_r.nbytes)

    # ---- leg 3 (auxiliary, BASELINE configs 4 and 5): whole-body tick alone on 65536 robots, and the full control
    #      tick (MPC update solve + whole-body tick fed with the MPC's first-step forces) on the MPC batch
    aux = {}
    if args.wbc_batch > 0:
        from qrw_b200 import BatchedWbcTick
        from qrw_b200 import fleet

        Bw = args.wbc_batch
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
        wt = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(wt, op=dist.ReduceOp.MAX)
        wbc_ms = float(wt.item()) / reps
        wbc_bytes = 1240  # 712 B in + 528 B out per tick through the batched-native API
        aux["wbc_tick"] = {"value": world * Bw / (wbc_ms * 1e-3), "unit": "ticks/s", "robots_per_gpu": Bw, "ms_per_step": wbc_ms,
                           "kernels_per_step": 3, "hbm_GBps_algorithmic": wbc_bytes * Bw / (wbc_ms * 1e-3) / 1e9,
                           "hbm_frac_of_measured": wbc_bytes * Bw / (wbc_ms * 1e-3) / 1e9 / measured_peaks()[0]["hbm_gbs"],
                           "workload": "BASELINE config 4: Solo12 kinematics/RNEA + InvKin + QPWBC + tau_ff, inputs resident in HBM"}
        # full control tick on the MPC batch: MPC update solve, then the whole-body tick with f_cmd = first-step forces
        tickm = BatchedWbcTick(B, device=local)
        margs = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in wl.wbc_tick_inputs(B, seed=5 + rank)]
        mpc.state_load(snap)
        for t in range(W):
            mpc.run(t + 1, *dev_in[t])
            margs[2] = fleet.first_step_forces(res).contiguous()
            tickm.compute(*margs)
        barrier()
        e0.record()
        for s_ in range(K):
            mpc.run(W + s_ + 1, *dev_in[W + s_])
            margs[2] = fleet.first_step_forces(res).contiguous()
            tickm.compute(*margs)
        e1.record()
        barrier()
        ft = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ft, op=dist.ReduceOp.MAX)
        aux["full_tick"] = {"value": world * B * K / (float(ft.item()) * 1e-3), "unit": "ticks/s", "robots_per_gpu": B,
                            "workload": "BASELINE config 5 shape: one MPC update solve + one whole-body tick per robot and step"}

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
        pt = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(pt, op=dist.ReduceOp.MAX)
        plan_ms = float(pt.item()) / 10
        plan_bytes = 6532  # 1,520 B read + 5,012 B written per robot and tick (xref 1,632 + fsteps 1,920 dominate)
        aux["planner_tick"] = {"value": world * Bw / (plan_ms * 1e-3), "unit": "ticks/s", "robots_per_gpu": Bw, "ms_per_step": plan_ms,
                               "kernels_per_step": 1, "hbm_GBps_algorithmic": plan_bytes * Bw / (plan_ms * 1e-3) / 1e9,
                               "hbm_frac_of_measured": plan_bytes * Bw / (plan_ms * 1e-3) / 1e9 / measured_peaks()[0]["hbm_gbs"],
                               "workload": "Gait + FootstepPlanner + FootTrajectoryGenerator + StatePlanner + feet commands, inputs resident in HBM"}
        del pl
        # closed control loop in HBM at the reference cadence (Controller.py:246: one MPC solve per k_mpc = 10 ticks):
        # planners -> MPC -> whole-body tick, one MPC period per CUDA-graph launch, inputs of the period staged on the device
        cw = wl.PlannerWorkload(B, 9 + rank, change_every=1000, yaw0=0.0, speed=0.5)
        cin = [[torch.from_numpy(a).to(dev) for a in cw.tick(k)] for k in range(10 * (2 + K))]
        ctl = BatchedController(B, device=local)
        for k in range(10):
            ctl.tick(k, *cin[k])
        stage = ctl.capture_period()

        def period(p_):
            for i in range(10):
                for dst, src in zip(stage, cin[10 * p_ + i]):
                    dst[i].copy_(src)
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
        ct = torch.tensor([period_ms.sum()], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ct, op=dist.ReduceOp.MAX)
        aux["control_loop"] = {"value": world * B * 10 * K / (float(ct.item()) * 1e-3), "unit": "control ticks/s", "robots_per_gpu": B,
                               "ms_per_mpc_period": float(ct.item()) / K, "period_ms": [round(float(x), 3) for x in period_ms],
                               "mpc_iters_mean_last_period": float(ctl.mpc.info()["iters"].mean()),
                               "kernels_per_period": 10 * 5 + 1, "periods": K,
                               "workload": "BASELINE config 5 at the reference cadence: 10 control ticks (planners + whole-body tick) and one "
                                           "MPC update solve per robot and period, everything resident in HBM, one CUDA graph per period"}
        ctl.close()

    # ---- roofline of the dominant kernel (the only kernel in a step)
    peaks, which = measured_peaks()
    kern_ms = float(step_ms.mean())
    achieved = BYTES_PER_SOLVE * B / (kern_ms * 1e-3) / 1e9
    roof = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": achieved / peaks["hbm_gbs"], "traffic": NCU_DRAM_BYTES_PER_SOLVE * B, "peak_source": which,
            "traffic_source": "ncu --set full, profiles/r01f_ncu_mpc_v2.txt, scaled by batch",
            "kernel": "mpc_v2_kernel", "algorithmic_bytes_per_solve": BYTES_PER_SOLVE}
    fp64 = None
    if hasattr(L, "qrw_fp64_peak"):
        import ctypes as C

        pk = C.c_double()
        if L.qrw_fp64_peak(local, C.byref(pk)) == 0:
            flop = float(timed_iters.mean()) * wl.FLOP_PER_ADMM_ITER + wl.FLOP_PER_SETUP
            ach = flop * B / (kern_ms * 1e-3) / 1e12
            fp64 = {"achieved": ach, "peak": pk.value, "unit": "TFLOP/s", "frac": ach / pk.value,
                    "flop_per_solve": flop, "peak_source": "DFMA microbenchmark in libqrw_b200 (qrw_fp64_peak)"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    cpu = None
    if not args.no_cpu_baseline:
        v, ms, cores, desc = cpu_leg(3, 1, args.cpu_sample)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_per_gpu": B, "n_steps": 16, "N_gait": 20, "history": "tick 0 cold (untimed) then update ticks, closed loop",
                   "l2": "256 MiB flush write between timed steps", "sharding": f"instances, {world} rank(s), all_gather of f0 only"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "roofline": roof, "fp64": fp64, "cpu_baseline": cpu, "clocks": clocks,
        "latency_ms": {"p50": float(np.percentile(step_ms, 50)), "p99": float(np.percentile(step_ms, 99)), "max": float(step_ms.max())},
        "admm_iters": {"mean": float(timed_iters.mean()), "p50": float(np.percentile(timed_iters, 50)),
                       "p99": float(np.percentile(timed_iters, 99)), "max": int(timed_iters.max())},
        "solved_frac": float((status == 1).mean()), "wall_s_device_leg": wall_dev,
    }
    line.update(aux)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
