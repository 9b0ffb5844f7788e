#!/usr/bin/env python
"""bench.py — IMC particle-segments/s of the jaybenne hot path on B200 (BASELINE.json metric).

Workload: synthetic 3D gray IMC on [0,1]^3, 256^3 cells in
32^3-cell meshblocks, periodic, rho = 1 + 0.1 sin sin sin, T0 = 1, kappa*rho*dx = 0.1,
sigma*dx = 0.5, dt = 4 dx / c, emission + feedback on, seed 349857.

One STEP = one pass of the hot path over one batch: reset the material fields, source ~1e9
thermal photons (SourcePhotons pass 1 + 2 and the census tally), then one full radiation
cycle (Fleck factor, emission sourcing, transport of every photon to census / absorption,
energy tallies, fluid update). Every step therefore does the same work (~1e9 photons,
~8.5 segments each); the swarm is cleared between steps. With N > 1 the meshblocks are
split over the ranks in Z-order chunks (strong scaling: the problem is the same 256^3 / 1e9)
and photons leaving a rank travel over NCCL inside jb_radiation_step.

  value   segments of all ranks / device time (CUDA events on the library's stream, max
          over ranks), inputs resident in HBM.
  e2e     same steps through the C ABI with HOST buffers: rho, sie copied from pinned host
          memory every step, energy_tally and internal_energy read back every step.
  roofline  transport kernel (k_transport): algorithmic bytes per launch
          84 B x (photons in + photons out) + 32 B x cells, over its own CUDA-event time.
  cpu_baseline  the CPU oracle (reference per-particle headers, OpenMP over photons, the
          reference's XorShift64 pool with one state per thread) on a 64^3-cell twin of the same workload on the box's host cores (rank 0, N = 1 only).

`--impl reference` times that CPU implementation alone (rank 0; other ranks exit 0).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C_LIGHT = 2.99792458e10
SEED = 349857


def workload_text(cells, nb, particles, stream=False, ddmc=False):
    from jaybenne_b200 import uniform_deck_text
    delta = 1.0 / cells
    t = uniform_deck_text([cells] * 3, [nb] * 3)
    t += f"""<parthenon/job>
problem_id = synthetic
<parthenon/time>
integrator = rk1
tlim = 1.0
<jaybenne>
num_particles = {particles}
dt = {(0.5 if stream else 4.0) * delta / C_LIGHT!r}
seed = {SEED}
do_emission = true
do_feedback = true
use_ddmc = {'true' if ddmc else 'false'}
tau_ddmc = 5.0
<mcblock>
opacity_model = constant
opacity_constant_value = {0.0 if stream else 0.1 / delta!r}
scattering_model = constant
scattering_constant_value = {0.0 if stream else (8.0 if ddmc else 0.5) / delta!r}
initial_density = 1.0
initial_temperature = 1.0
initial_radiation = thermal
"""
    return t


def density(x, y, z):
    import numpy as np
    return 1.0 + 0.1 * np.sin(2 * np.pi * x) * np.sin(2 * np.pi * y) * np.sin(2 * np.pi * z)


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU while the timed region runs (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def finish(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------
def cpu_arm(cells, nb, particles, steps, warmup, budget_s=300.0):
    """The reference's CPU implementation of the path (oracle) on the host cores.

    Returns (segments/s, ms/step, info). Each step is the same sequence as a GPU step on
    a bounded twin of the workload (same per-cell loading); if the first step shows that
    steps+warmup would not fit the time budget, the twin is shrunk once.
    """
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob
    from jaybenne_b200 import Mcblock, ParameterInput, capi
    ref_fast = os.path.join(ob.ORACLE_DIR, "_ref", "libjb_oracle_ref_fast.so")
    if os.path.exists(ref_fast):
        ob.FLAVOURS["ref_fast"] = ref_fast
        flavour, kind = "ref_fast", "reference"
    else:
        flavour, kind = "fast", "port"
    api = ob.oracle_api(flavour)
    cores = os.cpu_count() or 1

    def make(cells_, particles_):
        pin = ParameterInput(text=workload_text(cells_, min(nb, cells_), particles_))
        # the reference's generator: one XorShift64 pool state per host thread (Kokkos OpenMP backend)
        mb = Mcblock(pin, api=api, cfg_overrides={"rng_mode": capi.JB_RNG_XORSHIFT_THREADS})
        return mb

    def one_step(mb):
        mb.jb.clear_particles()
        mb.ProblemGenerator(density)           # fields + thermal sourcing + tally
        return mb.jb.RadiationStep(0.0, mb.cfg.dt)

    mb = make(cells, particles)
    t0 = time.perf_counter()
    st = one_step(mb)
    t_first = time.perf_counter() - t0
    done_warm = 1
    if t_first * (steps + warmup) > budget_s and cells > 32:
        cells, particles = cells // 2, particles // 8
        mb = make(cells, particles)
        done_warm = 0
    for _ in range(max(warmup - done_warm, 0)):
        one_step(mb)
    seg = 0
    t0 = time.perf_counter()
    for _ in range(steps):
        st = one_step(mb)
        seg += st["n_segments"]
    el = time.perf_counter() - t0
    info = {"kind": kind, "cores": cores, "flavour": api.flavour().decode(),
            "sample": f"{cells}^3-cell twin of the workload (same per-cell loading), {particles} photons, "
                      f"{steps} step(s): thermal sourcing + one cycle, OpenMP over photons, one XorShift64 pool state per thread as in the reference"}
    return seg / el, 1e3 * el / max(steps, 1), info


# ------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cells", type=int, default=256, help="cells per dimension (256 = the BASELINE config)")
    ap.add_argument("--block", type=int, default=32, help="cells per meshblock per dimension")
    ap.add_argument("--particles", type=int, default=0, help="photons per step (default 59.6 per cell -> 1e9 at 256^3)")
    ap.add_argument("--sort-interval", type=int, default=0)
    ap.add_argument("--gray-direct", action="store_true", help="sigma_a = rho*kappa instead of the reference's B_nu round trip")
    ap.add_argument("--stream", action="store_true",
                    help="C5-stream variant (SURVEY 8d): dt = dx/(2c), no opacity -> ~1.4 segments per photon, the "
                         "bandwidth-bound limit of the same kernels (not the BASELINE metric's config)")
    ap.add_argument("--ddmc", action="store_true",
                    help="hybrid variant: use_ddmc = true, tau_ddmc = 5 with sigma*dx = 8 (every cell optically thick -> "
                         "the DDMC kernel, transport_ddmc.cpp); not the BASELINE metric's config")
    ap.add_argument("--continuing", type=int, default=0, metavar="N",
                    help="extra report: after the timed steps, one thermal initialisation followed by N consecutive "
                         "cycles of the same simulation (emission + feedback; the population decays by ~1/3 per cycle, "
                         "holes accumulate and are squeezed by occupancy) -> config.continuing in the JSON line")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-cells", type=int, default=64)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cells, nb = args.cells, min(args.block, args.cells)
    particles = args.particles or int(round(1.0e9 * (cells / 256.0) ** 3))
    workload = (f"C5 synthetic 3D gray IMC: {cells}^3 cells, {nb}^3-cell meshblocks, periodic, {particles} thermal "
                f"photons sourced + one full cycle per step (dt = 4dx/c, kappa*dx = 0.1, sigma*dx = 0.5)")
    if args.stream:
        workload = workload.replace("C5 synthetic", "C5-stream synthetic").replace(
            "(dt = 4dx/c, kappa*dx = 0.1, sigma*dx = 0.5)", "(dt = dx/(2c), kappa = sigma = 0: free streaming)")
    if args.ddmc:
        workload = workload.replace("C5 synthetic", "C5-ddmc synthetic").replace("sigma*dx = 0.5)", "sigma*dx = 8, use_ddmc, tau_ddmc = 5)")
    config = {"workload": workload, "cells": cells, "meshblock": nb, "photons_per_step": particles,
              "parallelism": f"meshblocks over {world} rank(s), Z-order chunks", "l2": "inputs_exceed_l2",
              "gray_absorption": "direct" if args.gray_direct else "kirchhoff (reference arithmetic)",
              "sort_interval": args.sort_interval}

    # -------------------------------------------------------------------- reference arm
    if args.impl == "reference":
        if rank != 0:
            return 0
        cpu_cells = min(args.cpu_cells, cells)
        cpu_particles = int(round(particles * (cpu_cells / cells) ** 3))
        v, ms, info = cpu_arm(cpu_cells, nb, cpu_particles, args.steps, args.warmup)
        line = {"impl": "reference", "metric": "IMC particle-segments/sec", "value": v, "unit": "segments/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config,
                "cpu_baseline": {"value": v, "unit": "segments/s", "cores": info["cores"], "kind": info["kind"],
                                 "sample": info["sample"]},
                "e2e": {"value": v, "unit": "segments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    # ------------------------------------------------------------------------- our arm
    import numpy as np
    import torch
    import torch.distributed as dist
    from jaybenne_b200 import Mcblock, ParameterInput, capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: jaybenne_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    pin = ParameterInput(text=workload_text(cells, nb, particles, args.stream, args.ddmc))
    nblocks_total = (cells // nb) ** 3
    over = {"source_nblocks": nblocks_total, "sort_interval": args.sort_interval,
            "gray_mode": capi.JB_GRAY_DIRECT if args.gray_direct else capi.JB_GRAY_KIRCHHOFF}
    # thermal photons + one cycle's emission (particles / nblocks_total, sourcing.cpp:68-69) + slack for the
    # uneven spread of photons over ranks
    cap = int(particles * (1.0 + 1.0 / nblocks_total) / world * (1.02 if world == 1 else 1.25)) + 200000
    mb = Mcblock(pin, rank=rank, n_ranks=world, capacity=cap, cfg_overrides=over)
    jb = mb.jb
    if world > 1:
        idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            import ctypes as C
            raw = C.create_string_buffer(128)
            assert jb.api.comm_unique_id(raw) == 0, "ncclGetUniqueId failed"
            idbuf.copy_(torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8))
        dist.broadcast(idbuf, 0)
        jb.comm_init(bytes(idbuf.cpu().numpy().tobytes()))

    # material inputs: device-resident copies (value arm) and pinned host copies (e2e arm)
    n1 = nb
    ncell = n1 ** 3
    gids = jb.local_gids
    host_rho = torch.empty((len(gids), ncell), dtype=torch.float64, pin_memory=True)
    host_sie = torch.empty((len(gids), ncell), dtype=torch.float64, pin_memory=True)
    host_tally = torch.empty((len(gids), ncell), dtype=torch.float64, pin_memory=True)
    host_u = torch.empty((len(gids), ncell), dtype=torch.float64, pin_memory=True)
    cv = mb.cfg.eos_cv
    for q, gid in enumerate(gids):
        xc = mb.mesh.cell_centers(gid)
        rho = density(xc[0][None, None, :], xc[1][None, :, None], xc[2][:, None, None])
        host_rho[q] = torch.from_numpy(np.ascontiguousarray(rho).ravel())
        host_sie[q] = cv * mb.tt0
    dev_rho, dev_sie = host_rho.cuda(), host_sie.cuda()
    dt = mb.cfg.dt

    def step(src_rho, src_sie, readback):
        jb.clear_particles()
        jb.set_fields_ptr("density", gids, [src_rho[q].data_ptr() for q in range(len(gids))])
        jb.set_fields_ptr("sie", gids, [src_sie[q].data_ptr() for q in range(len(gids))])
        jb.InitializeRadiation(True)
        jb.PostInitialization()
        st = jb.RadiationStep(0.0, dt)
        if readback:
            jb.get_fields_ptr("energy_tally", gids, [host_tally[q].data_ptr() for q in range(len(gids))])
            jb.get_fields_ptr("internal_energy", gids, [host_u[q].data_ptr() for q in range(len(gids))])
        return st

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_sum(x):
        if world == 1:
            return x
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        return float(t.item())

    def reduce_max(x):
        if world == 1:
            return x
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(args.warmup):
        step(dev_rho, dev_sie, False)

    # ---- timed: device-resident inputs
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    launches0 = jb.kernel_launches()
    jb.timer_start()
    seg = 0
    tr_ms = 0.0
    tr_bytes = 0.0
    cyc_bytes = 0.0
    last = None
    for _ in range(args.steps):
        st = step(dev_rho, dev_sie, False)
        seg += st["n_segments"]
        tr_ms += st["ms_transport"]
        n_in = st["n_start"] + st["n_sourced"] + st["n_received"]
        n_out = st["n_end"] + st["n_sent"]
        tr_bytes += 84.0 * (n_in + n_out) + 32.0 * len(gids) * ncell
        cyc_bytes += 84.0 * (n_in + n_out) + 84.0 * (st["n_start"] + st["n_sourced"]) + 40.0 * len(gids) * ncell
        last = st
    ms = jb.timer_stop()
    launches = jb.kernel_launches() - launches0
    barrier()
    clocks = sampler.finish()
    ms_max = reduce_max(ms)
    seg_all = reduce_sum(seg)
    value = seg_all / (ms_max * 1e-3)
    tr_ms_max = reduce_max(tr_ms)

    # ---- timed: end to end through the C ABI with host buffers
    e2e = None
    if not args.no_e2e:
        step(host_rho, host_sie, True)  # warm the host path once
        barrier()
        seg2 = 0
        t0 = time.perf_counter()
        for _ in range(args.steps):
            st = step(host_rho, host_sie, True)
            seg2 += st["n_segments"]
        torch.cuda.synchronize()
        el = time.perf_counter() - t0
        el_max = reduce_max(el)
        seg2_all = reduce_sum(seg2)
        nb_local = reduce_sum(len(gids))
        e2e = {"value": seg2_all / el_max, "unit": "segments/s",
               "h2d_bytes_per_step": int(2 * 8 * ncell * nb_local), "d2h_bytes_per_step": int(2 * 8 * ncell * nb_local),
               "result_checksum": float(host_tally.sum().item())}

    # ---- optional: a continuing simulation (steady-state behaviour of compaction / slot order)
    continuing = None
    if args.continuing > 0:
        jb.clear_particles()
        jb.set_fields_ptr("density", gids, [dev_rho[q].data_ptr() for q in range(len(gids))])
        jb.set_fields_ptr("sie", gids, [dev_sie[q].data_ptr() for q in range(len(gids))])
        jb.InitializeRadiation(True)
        jb.PostInitialization()
        barrier()
        jb.timer_start()
        cyc = []
        t_now = 0.0
        for _ in range(args.continuing):
            st = jb.RadiationStep(t_now, dt)
            jb.UpdateDerivedMaterial()
            t_now += dt
            cyc.append({"n_start": st["n_start"], "n_segments": st["n_segments"], "ms_total": st["ms_total"],
                        "ms_transport": st["ms_transport"]})
        cms = reduce_max(jb.timer_stop())
        cseg = reduce_sum(sum(c["n_segments"] for c in cyc))
        continuing = {"cycles": args.continuing, "value": cseg / (cms * 1e-3), "unit": "segments/s", "ms": cms,
                      "per_cycle_rank0": cyc}

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = tr_bytes / (tr_ms * 1e-3) / 1e9 if tr_ms > 0 else 0.0
    # DRAM traffic of the kernel: measured once with ncu on the 64^3 twin (profiles/r01_traffic.json) as a ratio to
    # the algorithmic bytes of that launch; reported here scaled to this run's launch (None if the file is absent)
    traffic, traffic_src, ncu_pipe = None, None, {}
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
        if not args.stream and not args.ddmc:
            traffic = tj["ratio"] * tr_bytes / max(args.steps, 1)
            traffic_src = f"ncu at {tj['cells']}^3: dram bytes / algorithmic bytes = {tj['ratio']}, scaled to this launch"
            ncu_pipe = {"fp64_pipe_pct": tj.get("fp64_pipe_pct"), "issue_active_pct": tj.get("issue_active_pct")}
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "k_transport<ddmc>" if args.ddmc else "k_transport_imc", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "ncu": ncu_pipe,
                "note": "8.3 segments per photon make this kernel FP64-issue bound (~220 FP64 ops per segment against "
                        "~17 algorithmic bytes); bench.py --stream measures the same kernel in the bandwidth-bound regime",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                "bytes_per_launch": tr_bytes / max(args.steps, 1), "ms_per_launch": tr_ms / max(args.steps, 1),
                "bytes_per_segment": tr_bytes / max(seg, 1),
                "whole_cycle_GBps": cyc_bytes / (ms * 1e-3) / 1e9,
                "segments_per_photon": seg / max(args.steps * (last["n_start"] + last["n_sourced"]), 1)}

    # second roofline (SURVEY 8d honesty note): FP64 pipe. Peak = DFMA chain kernel measured now on this GPU;
    # flops per segment = DFMA x2 + DMUL + DADD of the kernel's main body, counted by ncu (profiles/r01_traffic.json)
    fp64 = None
    try:
        fpeak = jb.measure_fp64_peak()
        fps = float(tj.get("fp64_flops_per_segment", 317.0)) if not args.stream else 317.0
        fach = (seg / (tr_ms * 1e-3)) * fps / 1e12 if tr_ms > 0 else 0.0
        fp64 = {"bound": "fp64", "peak": fpeak, "unit": "TFLOP/s", "peak_source": "k_fp64_chain measured in this run",
                "flops_per_segment": fps, "achieved": fach, "frac": fach / fpeak if fpeak > 0 else None,
                "pipe_active_pct_ncu": ncu_pipe.get("fp64_pipe_pct")}
    except Exception as exc:  # noqa: BLE001
        fp64 = {"error": str(exc)}
    roofline["fp64"] = fp64

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_cells = min(args.cpu_cells, cells)
        cpu_particles = int(round(particles * (cpu_cells / cells) ** 3))
        v, cms, info = cpu_arm(cpu_cells, nb, cpu_particles, 1, 1)
        cpu = {"value": v, "unit": "segments/s", "cores": info["cores"], "kind": info["kind"], "sample": info["sample"],
               "ms_per_step": cms}

    if rank == 0:
        line = {"metric": "IMC particle-segments/sec", "value": value, "unit": "segments/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / max(args.steps, 1),
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config, "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
                "roofline": roofline, "cpu_baseline": cpu,
                "transport_ms_per_step": tr_ms_max / max(args.steps, 1), "continuing": continuing,
                "last_step": {k: last[k] for k in ("n_start", "n_sourced", "n_absorbed", "n_end", "n_sent", "n_received",
                                                    "n_segments", "n_scatters", "n_block_crossings", "n_iterations")}}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
