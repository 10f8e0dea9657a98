#!/usr/bin/env python
"""bench.py -- direct-N^2 G-interactions/s (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--bodies B]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one Velocity-Verlet step (store-acc + drift, N^2 force evaluation, kick) =
Integrator::integrate of the reference (src/cuda/integrator.cu:224-238) = N^2 pair interactions.
  N = 1 : configs[1] of BASELINE.json -- 1 Mi bodies, Plummer, fp32, one B200.
  N > 1 : configs[2] -- 4 Mi bodies target-sharded over N GPUs; the all-gather of packed positions is
          fused into our own kernels over NVLink peer memory (NB_MGPU_EXCHANGE=nccl selects the NCCL
          all-gather overlapped with the local-source pass instead).  Strong scaling: total work fixed.
`value` times K steps with state resident in HBM; `e2e` times the host-buffer C-ABI call
(nb_host_step_direct / ShardedDirectSystem.host_update): pinned host state in, one step, host state
out, copies inside the timed region.  The CPU legs (cpu_baseline, --impl reference) run the oracle
port of the reference's CPU loop (examples/example_force_methods.cpp:44-66) on a bounded sample.

After the timed regions (never inside them) the line also carries
  parity        every N: max relative error of >= 256 sampled targets per rank of the system that was just timed,
                against the fp64 oracle over ALL sources (all-reduced MAX); bound 1e-5;
  exchange_nccl N > 1: the same step with the NCCL all_gather_into_tensor exchange, 3 steps, next to the peer push;
  t1_4mi        N = 1: the 4 Mi-body workload of configs[2] on ONE GPU, so that strong-scaling efficiency can be
                stated as T1 / (G * T_G);
  paths         N = 1: configs[3] and [4] -- Barnes-Hut theta = 0.5 at 1 Mi and 16 Mi bodies, spatial hash at 10 M,
                the radix sort at 16 Mi x 30 bits -- per stage: CUDA-event ms, algorithmic bytes, achieved GB/s and
                the fraction of the measured HBM bandwidth (tools/bench_paths.py; the reference benchmarks all
                three methods from one program too, benchmarks/benchmark_main.cpp:170-233).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

if "reference" in sys.argv:
    # torchrun exports OMP_NUM_THREADS=1; the CPU arm uses every host thread, and libgomp reads the
    # variable when it is first loaded, so fix it before anything can pull OpenMP in.
    os.environ["OMP_NUM_THREADS"] = str(len(os.sched_getaffinity(0)))

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_INTERACTION = 20.0
FP32_PEAK_DATASHEET_TF = 80.0    # B200 datasheet, the denominator BASELINE.json's target names
G_CONST, SOFTENING, DT = 1.0, 0.01, 1e-3


_JSON_FD = None


def quiet_stdout():
    """Libraries below us print to fd 1 (NCCL's version banner under NCCL_DEBUG): send fd 1 to stderr while
    the bench runs and keep the real stdout for the one JSON line."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_JSON_FD if _JSON_FD is not None else 1, data)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bodies", type=int, default=0, help="override the body count (parity/debug only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-paths", action="store_true", help="skip the Barnes-Hut / spatial-hash / sort / 4 Mi legs")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--path", default="direct", choices=["direct", "bh"],
                    help="direct: the headline direct-N^2 metric (default); bh: the Barnes-Hut step of configs[3], 16 Mi "
                         "bodies, theta 0.5, traversal target-sharded over replicated state at N > 1")
    ap.add_argument("--no-reference-gpu", action="store_true")
    return ap.parse_args()


def workload(n_gpus: int, bodies: int):
    if bodies:
        return bodies, f"direct_n2_{bodies}_plummer"
    if n_gpus == 1:
        return 1 << 20, "direct_n2_1Mi_plummer_1gpu"
    return 1 << 22, "direct_n2_4Mi_plummer_target_sharded"


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 8:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2])); pw.append(float(p[3]))
            except ValueError:
                continue
            for nm, v in zip(names, p[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_mhz_min": float(min(sm)), "sm_mhz_mean": float(np.mean(sm)),
                "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)), "power_w_mean": float(np.mean(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_sample_targets(n: int, per_thread: int) -> int:
    """Targets in the bounded CPU sample: per_thread targets for every host thread (one thread does
    ~0.35 G-interactions/s, so 4096 targets x 1 Mi sources per thread is ~12 s of CPU work)."""
    return min(n, per_thread * len(os.sched_getaffinity(0))) if n >= (1 << 20) else min(n, 4096)


def cpu_sample_rate(state, n_targets: int, reps: int = 1):
    """Oracle port of the reference CPU loop on a bounded sample: the first n_targets targets of the
    workload against ALL sources.  Returns (G-interactions/s, threads, seconds)."""
    from oracle import bindings as orc
    orc.set_num_threads(len(os.sched_getaffinity(0)))  # all host threads, whatever OMP_NUM_THREADS says
    n = state["pos_x"].size
    n_targets = min(n_targets, n)
    t0 = time.perf_counter()
    for _ in range(reps):
        orc.direct_f32(state["pos_x"], state["pos_y"], state["pos_z"], state["mass"], G_CONST,
                       float(np.float32(SOFTENING) ** 2), 0, n_targets)
    dt = (time.perf_counter() - t0) / reps
    return n_targets * float(n) / dt / 1e9, orc.num_threads(), dt


def parity_check(system, state, n, begin, count, rank, world, dist, torch):
    """Accelerations of the system that was just timed against the fp64 oracle: the CURRENT positions of all bodies
    (gathered from the ranks), one more force evaluation (collective), >= 256 targets of this rank's slice against
    ALL sources.  Returns the all-reduced maximum of |a - a_ref| / |a_ref|."""
    from oracle import bindings as orc
    keys = ("pos_x", "pos_y", "pos_z")
    local = {k: system.d.arrays[k].detach().cpu().numpy().copy() for k in keys}
    if world > 1:
        parts = [None] * world
        dist.all_gather_object(parts, local)
        full = {k: np.concatenate([p[k] for p in parts]) for k in keys}
    else:
        full = local
    assert full["pos_x"].size == n
    system.compute_forces()
    torch.cuda.synchronize()
    acc = {k: system.d.arrays[k].detach().cpu().numpy() for k in ("acc_x", "acc_y", "acc_z")}
    m = max(256, 1)
    idx_local = np.unique(np.concatenate([np.linspace(0, count - 1, m).astype(np.int64), [0, count - 1]]))
    idx = (idx_local + begin).astype(np.int32)
    orc.set_num_threads(max(1, len(os.sched_getaffinity(0)) // world))
    ref = orc.direct_f64(full["pos_x"], full["pos_y"], full["pos_z"], state["mass"], G_CONST,
                         float(np.float32(SOFTENING) ** 2), idx)
    err = np.sqrt(sum((acc[k][idx_local].astype(np.float64) - r) ** 2 for k, r in zip(("acc_x", "acc_y", "acc_z"), ref)))
    rel = float((err / np.sqrt(sum(r ** 2 for r in ref))).max())
    if world > 1:
        t = torch.tensor([rel], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        rel = float(t.item())
    return {"max_rel_err": rel, "bound": 1e-5, "ok": bool(rel <= 1e-5), "targets_per_rank": int(idx.size),
            "sources": n, "oracle": "oracle/direct.c orc_direct_f64 (fp64, all sources), on the positions the timed "
                                    "steps left behind"}


def partial_slots_bytes(ctx, L):
    """Bytes the library's scratch holds right now (packed sources + per-chunk partials dominate after a direct run)."""
    try:
        return int(L.nb_scratch_bytes(ctx.handle))
    except Exception:
        return None


def direct_traffic(n, world, scratch_bytes):
    """roofline.traffic: dram__bytes_read.sum + dram__bytes_write.sum of ONE force-kernel launch from the ncu --set
    full capture of this command, kept in profiles/direct_traffic.json by tools/ncu_summary.py (null when no capture
    of the current build exists); scratch_bytes is what the library held during THIS run."""
    out = {"traffic": None, "traffic_unit": "bytes per launch (compute-bound kernel)", "scratch_bytes": scratch_bytes}
    try:
        with open(os.path.join(ROOT, "profiles", "direct_traffic.json")) as f:
            rec = json.load(f)
        if rec.get("bodies") == n and world == 1:
            out["traffic"] = rec["dram_bytes_per_launch"]
            out["traffic_source"] = rec.get("source")
    except Exception:
        pass
    return out


def run_bh_path(args):
    """--path bh: configs[3] as a timed step -- store-acc + drift, tree rebuild (bbox, Morton keys, radix sort, tree +
    centres of mass + groups), group traversal, kick -- for 16 Mi Plummer bodies, theta 0.5.  At N > 1 every rank holds
    all bodies and rebuilds the tree, traverses only the groups dealt to it and pushes the accelerations to every
    replica over NVLink peer memory (nbody_b200/dist.py ShardedBarnesHutSystem).  Strong scaling (total work fixed).
    Parity: the replicas after the timed steps against the single-GPU step on the same inputs, bit for bit."""
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    import nbody_b200 as nb
    from nbody_b200 import ics
    from nbody_b200.dist import ShardedBarnesHutSystem
    n = args.bodies or (1 << 24)
    theta = 0.5
    st = ics.plummer(n, seed=42)
    ctx = nb.get_context(local_rank)
    sysm = ShardedBarnesHutSystem(st, rank, world, G_CONST, SOFTENING, theta)
    sysm.compute_forces()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(args.warmup):
        sysm.update(DT)
    barrier()
    launches0 = ctx.launch_count
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        sysm.update(DT)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    launches = ctx.launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None
    got = sysm.arrays()
    sysm.close()
    parity = None
    if not args.no_parity:
        single = nb.ParticleSystem()
        single.config_ = nb.SimulationConfig(particle_count=n, softening=SOFTENING, dt=DT, G=G_CONST,
                                             barnes_hut_theta=theta, force_method=nb.ForceMethod.BARNES_HUT)
        single.setState(nb.SimulationState(n, 0.0, DT, G_CONST, SOFTENING, nb.ForceMethod.BARNES_HUT,
                                           {k: st[k] for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "mass")}))
        for _ in range(args.warmup + args.steps):
            single.update(DT)
        torch.cuda.synchronize()
        want = single.getDeviceData().numpy()
        same = all(np.array_equal(got[k], want[k]) for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z",
                                                             "acc_x", "acc_y", "acc_z"))
        ok = max_over_ranks(0.0 if same else 1.0) == 0.0
        parity = {"replicas_bit_identical_to_single_gpu_step": bool(ok), "ok": bool(ok),
                  "steps_compared": args.warmup + args.steps}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    emit({"metric": "barnes_hut_mbody_steps_per_s", "value": n / (ms * 1e-3) / 1e6, "unit": "M body-steps/s",
          "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
          "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
          "config": {"workload": f"barnes_hut_theta0.5_{n}_plummer", "bodies": n, "theta": theta, "softening": SOFTENING,
                     "dt": DT, "parallelism": f"replicated_state_traversal_sharded_x{world}" if world > 1 else "single_gpu",
                     "exchange": "accelerations pushed to every replica from the traversal kernel's epilogue (peer memory)"
                     if world > 1 else "none"},
          "gpu_launches": int(launches), "clocks": clocks, "parity": parity})


def run_reference_arm(args):
    """--impl reference: the reference has no CPU force path of its own; the timed
    CPU implementation is the oracle port of its example loop, all host threads, bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from nbody_b200 import ics
    n, wl = workload(args.gpus, args.bodies)
    st = ics.plummer(n, seed=42)
    sample = cpu_sample_targets(n, 1024)  # ~3 s per step: the whole run stays within a few minutes
    for _ in range(max(args.warmup, 0)):
        cpu_sample_rate(st, max(sample // 8, 256))
    times = []
    for _ in range(max(args.steps, 1)):
        rate, threads, sec = cpu_sample_rate(st, sample)
        times.append(sec)
    sec = float(np.mean(times))
    value = sample * float(n) / sec / 1e9
    line = {
        "impl": "reference", "metric": "direct_n2_g_interactions_per_s", "value": value,
        "unit": "G-interactions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl, "bodies": n, "ic": "plummer_henon_seed42", "softening": SOFTENING, "dt": DT,
                   "step": "sample of one force evaluation"},
        "cpu_baseline": {"value": value, "unit": "G-interactions/s", "cores": threads, "kind": "port",
                         "sample": f"first {sample} targets x all {n} sources per step "
                                   f"({sample * n / 1e9:.1f} G pair interactions), OpenMP over targets"},
        "e2e": {"value": value, "unit": "G-interactions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def main():
    args = parse()
    quiet_stdout()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    if args.path == "bh":
        run_bh_path(args)
        return

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("launch N>1 with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    from nbody_b200.dist import ShardedDirectSystem, shard_layout, slice_state

    n, wl = workload(args.gpus, args.bodies)
    state = ics.plummer(n, seed=42)
    eps2 = float(np.float32(SOFTENING) ** 2)
    ctx = nb.get_context(local_rank)
    L = _lib.lib()

    begin, count, _ = shard_layout(n, rank, world)
    system = ShardedDirectSystem(slice_state(state, begin, count), n, rank, world, G_CONST, SOFTENING)
    system.compute_forces()  # a(0), ParticleSystem::initialize (particle_system.cpp:91)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident timing ---------------------------------------------------------------
    for _ in range(args.warmup):
        system.update(DT)
    barrier()
    L.nb_profile_enable(ctx.handle, 1)
    launches0 = ctx.launch_count
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        system.update(DT)
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = ctx.launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None
    kms, kl = C.c_double(0), C.c_uint64(0)
    L.nb_profile_read(ctx.handle, C.byref(kms), C.byref(kl))
    L.nb_profile_enable(ctx.handle, 0)
    ms_per_step = ms_total / args.steps
    interactions_per_step = float(n) * float(n)
    value = interactions_per_step / (ms_per_step * 1e-3) / 1e9

    # ---- end-to-end through the host-buffer call ------------------------------------------------
    for _ in range(min(args.warmup, 2)):
        system.host_update(DT, 1)
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        system.host_update(DT, 1)
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(max(e0.elapsed_time(e1), 0.0)) / args.steps
    e2e_wall = max_over_ranks((time.perf_counter() - t0) * 1e3) / args.steps
    e2e_ms = max(e2e_ms, e2e_wall)  # the host clock also sees the final D2H and sync
    e2e_value = interactions_per_step / (e2e_ms * 1e-3) / 1e9
    h2d = 10 * 4 * count
    d2h = 12 * 4 * count

    # ---- parity of the system that was just timed (after timing): fp64 oracle on sampled targets ------
    parity = None
    if not args.no_parity:
        parity = parity_check(system, state, n, begin, count, rank, world, dist if world > 1 else None, torch)

    exchange = system.exchange
    partial_bytes = partial_slots_bytes(ctx, L)
    system.close()

    # ---- N > 1: the collective north_star names (NCCL all-gather overlapped with the local pass) ------
    exchange_nccl = None
    if world > 1 and exchange != "nccl":
        sys_n = ShardedDirectSystem(slice_state(state, begin, count), n, rank, world, G_CONST, SOFTENING, exchange="nccl")
        sys_n.compute_forces()
        sys_n.update(DT)
        barrier()
        k3 = 3
        e0.record()
        for _ in range(k3):
            sys_n.update(DT)
        e1.record()
        barrier()
        ms_n = max_over_ranks(e0.elapsed_time(e1)) / k3
        sys_n.host_update(DT, 1)
        barrier()
        t0 = time.perf_counter()
        for _ in range(k3):
            sys_n.host_update(DT, 1)
        barrier()
        e2e_n = max_over_ranks((time.perf_counter() - t0) * 1e3) / k3
        sys_n.close()
        exchange_nccl = {"ms_per_step": ms_n, "value": interactions_per_step / (ms_n * 1e-3) / 1e9,
                         "e2e_ms_per_step": e2e_n, "e2e_value": interactions_per_step / (e2e_n * 1e-3) / 1e9,
                         "unit": "G-interactions/s", "steps": k3,
                         "what": "same workload, exchange = NCCL all_gather_into_tensor of the packed slices on NCCL's "
                                 "stream, overlapped with the local-source pass (nbody_b200/dist.py)",
                         "peer_over_nccl": ms_n / ms_per_step}

    # ---- the same bodies through the general pair loop -------------------------------------------
    # The workload's masses are all 1/N (as in every initial condition of the reference), so every full
    # source tile takes the one-mass pair loop.  For the record, the same step with the last
    # mantissa bit of every 512th mass flipped, which sends every tile through the general loop.
    mixed = None
    if world == 1:
        st2 = {k: v.copy() for k, v in state.items()}
        st2["mass"].view(np.uint32)[::512] ^= 1
        sys2 = ShardedDirectSystem(st2, n, 0, 1, G_CONST, SOFTENING)
        sys2.compute_forces()
        sys2.update(DT)
        torch.cuda.synchronize()
        k2 = max(1, min(args.steps, 3))
        e0.record()
        for _ in range(k2):
            sys2.update(DT)
        e1.record()
        torch.cuda.synchronize()
        ms2 = e0.elapsed_time(e1) / k2
        sys2.close()
        mixed = {"value": interactions_per_step / (ms2 * 1e-3) / 1e9, "unit": "G-interactions/s", "ms_per_step": ms2,
                 "steps": k2, "what": "same bodies, one mass in every 512 differs in its last bit: general pair loop "
                                      "(3 packed multiplies per pair instead of 2) for every source tile"}
    t1_4mi, paths, probe = None, None, None
    if world == 1 and not args.no_paths and not args.bodies:
        # FP32 pipe rates of THIS device, measured now (replaces constants copied from an earlier capture)
        probe = {}
        for name, packed in (("ffma", 0), ("ffma2", 1)):
            tf, pms = C.c_double(0), C.c_double(0)
            _lib.check(L.nb_probe_fp32(ctx.handle, packed, C.byref(tf), C.byref(pms), None))
            probe[name] = {"tflops": tf.value, "ms": pms.value}
        # configs[2]'s workload on one GPU: T1 of the strong-scaling efficiency
        n4 = 1 << 22
        st4 = ics.plummer(n4, seed=42)
        sys4 = ShardedDirectSystem(st4, n4, 0, 1, G_CONST, SOFTENING)
        sys4.update(DT)
        torch.cuda.synchronize()
        k4 = 2
        e0.record()
        for _ in range(k4):
            sys4.update(DT)
        e1.record()
        torch.cuda.synchronize()
        ms4 = e0.elapsed_time(e1) / k4
        sys4.close()
        del sys4, st4
        t1_4mi = {"ms_per_step": ms4, "value": float(n4) * n4 / (ms4 * 1e-3) / 1e9, "unit": "G-interactions/s",
                  "steps": k4, "warmup": 1, "bodies": n4, "workload": "direct_n2_4Mi_plummer on ONE GPU",
                  "what": "T1 for strong-scaling efficiency T1 / (G * T_G) of the N > 1 runs (SURVEY.md 8(d) config 3)"}
        scratch_direct = int(L.nb_scratch_bytes(ctx.handle))
        # configs[3], configs[4] and the sort, per stage (tools/bench_paths.py)
        from tools import bench_paths as bp
        paths = {"bh_1mi": bp.bench_bh(1 << 20), "bh_16mi": bp.bench_bh(1 << 24), "sh_10m": bp.bench_sh(10_000_000),
                 "sort_16mi": bp.bench_sort(1 << 24, 30), "sort_10m_21bit": bp.bench_sort(10_000_000, 21),
                 "scratch_bytes_after_direct_legs": scratch_direct,
                 "scratch_bytes_after_all_legs": int(L.nb_scratch_bytes(ctx.handle))}
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return

    # ---- roofline of the dominant kernel (FP32 CUDA cores; neither HBM nor tensor bound) ---------
    peaks = measured_peaks()
    sm_max = float(peaks.get("sm_max_mhz", 1965.0))
    clock_peak_tf = 148 * 128 * 2 * sm_max * 1e6 / 1e12
    kernel_ms = kms.value / max(kl.value, 1)                  # per force-kernel launch (CUDA events)
    pairs_per_launch = float(count) * float(n)
    if world > 1:
        kernel_ms = kms.value / args.steps                    # local + remote pass of one step
    achieved_tf = pairs_per_launch * FLOP_PER_INTERACTION / (kernel_ms * 1e-3) / 1e12 if kernel_ms > 0 else None
    roofline = {
        "bound": "fp32", "kernel": "nb::direct_force_kernel<256,2,1024>",
        "achieved": achieved_tf, "peak": FP32_PEAK_DATASHEET_TF, "unit": "TFLOP/s",
        "frac": (achieved_tf / FP32_PEAK_DATASHEET_TF) if achieved_tf else None,
        "peak_source": "B200 datasheet FP32 (BASELINE.json target denominator); MEASURED_PEAKS.json holds no FP32 figure",
        "peak_clock_derived": clock_peak_tf, "frac_of_clock_peak": (achieved_tf / clock_peak_tf) if achieved_tf else None,
        # scalar FFMA / packed FFMA2 rates measured by THIS run (nb_probe_fp32), null when the legs are skipped
        "peak_measured_ffma": probe["ffma"]["tflops"] if probe else None,
        "peak_measured_ffma2": probe["ffma2"]["tflops"] if probe else None,
        "frac_of_measured_ffma2": (achieved_tf / probe["ffma2"]["tflops"]) if (achieved_tf and probe) else None,
        "flop_per_interaction": FLOP_PER_INTERACTION, "kernel_ms": kernel_ms,
        "kernel_share_of_step": (kms.value / args.steps) / ms_per_step if ms_per_step > 0 else None,
        **direct_traffic(n, world, partial_bytes),
    }

    cpu = None
    if not args.no_cpu_baseline and world == 1:   # rank 0 at N = 1 only
        sample = cpu_sample_targets(n, 4096)
        rate, threads, sec = cpu_sample_rate(state, sample)
        cpu = {"value": rate, "unit": "G-interactions/s", "cores": threads, "kind": "port",
               "sample": f"first {sample} targets x all {n} sources of the same workload, {sec:.1f} s, "
                         "oracle port of examples/example_force_methods.cpp:44-66, OpenMP over targets"}

    ref_gpu = None
    if not args.no_reference_gpu and world == 1:
        try:
            from oracle import ref_bindings as ref
            if ref.available():
                nr = min(n, 1 << 18)   # the reference kernel is slow; a 256 Ki-body evaluation is enough
                sub = {k: np.ascontiguousarray(v[:nr]) for k, v in state.items()}
                ms = ref.forces(sub, 0, G_CONST, SOFTENING, iters=2)
                ref_gpu = {"value": float(nr) * nr / (ms * 1e-3) / 1e9, "unit": "G-interactions/s",
                           "what": "reference computeForcesDirectKernel (unmodified source, sm_100) on this B200",
                           "bodies": nr, "ms": ms}
        except Exception as e:  # the reference build is optional evidence, never the product path
            ref_gpu = {"unavailable": str(e)[:200]}

    line = {
        "metric": "direct_n2_g_interactions_per_s", "value": value, "unit": "G-interactions/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "steps_per_s": 1e3 / ms_per_step, "higher_is_better": True,
        "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl, "bodies": n, "ic": "plummer_henon_seed42", "softening": SOFTENING, "dt": DT,
                   "G": G_CONST, "parallelism": f"target_sharded_x{world}" if world > 1 else "single_gpu",
                   "exchange": ("peer-memory push fused into the pack kernel, flag wait between the local and remote pass (no NCCL in the step)"
                                if exchange == "peer" else "NCCL all_gather_into_tensor overlapped with the local pass")
                   if world > 1 else "none",
                   "masses": "all equal (1/N): source tiles of one mass take the one-mass pair loop, decided per tile "
                             "on the device from the data; see mixed_mass_path for the general loop",
                   "cache": "positions (16 MiB+ packed) are rewritten every step; compute bound, L2 state irrelevant"},
        "tflops_20flop": value * 1e9 * FLOP_PER_INTERACTION / 1e12,
        "pct_fp32_peak_datasheet": 100.0 * value * 1e9 * FLOP_PER_INTERACTION / 1e12 / (FP32_PEAK_DATASHEET_TF * args.gpus),
        "e2e": {"value": e2e_value, "unit": "G-interactions/s", "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        "reference_gpu": ref_gpu, "mixed_mass_path": mixed, "parity": parity, "exchange_nccl": exchange_nccl,
        "t1_4mi": t1_4mi, "paths": paths,
    }
    emit(line)


if __name__ == "__main__":
    main()
