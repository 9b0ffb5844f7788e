#!/usr/bin/env python
"""bench.py — IMC particle-segments/s of the jaybenne hot path on B200 (BASELINE.json metric).

Workload: synthetic 3D gray IMC on [0,1]^3, 256^3 cells in
32^3-cell meshblocks, periodic, rho = 1 + 0.1 sin sin sin, T0 = 1, kappa*rho*dx = 0.1,
sigma*dx = 0.5, dt = 4 dx / c, emission + feedback on, seed 349857.

One STEP = one pass of the hot path over one batch: reset the material fields, source ~1e9
thermal photons (SourcePhotons pass 1 + 2 and the census tally), then one full radiation
cycle (Fleck factor, emission sourcing, transport of every photon to census / absorption,
energy tallies, fluid update). The per-photon streams are restarted every step
(jb_reset_streams), so every step is the SAME batch (~1e9 photons, ~8.3 segments each) and
`invariants` (counts and energy sums all-reduced over the ranks) must be equal for every N,
K and W. With N > 1 the meshblocks are split over the ranks in Z-order chunks (strong
scaling: the problem is the same 256^3 / 1e9) and photons leaving a rank travel over NVLink
inside jb_radiation_step.

  value   segments of all ranks / device time (CUDA events on the library's stream, max
          over ranks), inputs resident in HBM.
  e2e     same steps through the C ABI with HOST buffers: rho, sie copied from pinned host
          memory every step, energy_tally and internal_energy read back every step.
  roofline  the transport kernel: algorithmic bytes per launch = 84 B x (photons in +
          photons out) + 32 B x cells, over its own CUDA-event time; `whole_cycle_GBps` uses
          SURVEY 8d's per-cycle bytes (... + 84 B x sourced + 40 B x cells) over the step.
          roofline.fp64 = the FP64 pipe (the bound that actually applies at 8.3 segments
          per photon), against a DFMA-chain peak measured in the same run.
  variants  timed in the same run, each a few seconds: `continuing` (8 consecutive cycles of one
          simulation), `continuing_sorted` (same, pool re-sorted by cell every cycle), `ddmc`
          (128^3 optically thick, k_transport_hpool) and `stream` (free streaming).
  timeline  device-timed phases of the last timed step (jb_step_timeline): rank 0's list, the sums
          per phase as max / min over the ranks, and which swarm exchange ran (direct peer
          stores over NVLink, or NCCL send / recv per iteration).
  cpu_baseline  the CPU oracle (reference per-particle headers, OpenMP over the photons of
          all blocks, the reference's XorShift64 pool with one state per thread) on a
          64^3-cell twin of the same workload on the box's host cores (rank 0, N = 1 only).

`--impl reference` times that CPU implementation alone (rank 0; other ranks exit 0); its
`config` names the twin it ran and its `cores` is what OpenMP reports after this script set
the thread count itself.
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

    Returns (segments/s, ms/step, info). Each step is the same sequence as a GPU step on a bounded twin of the
    workload (same per-cell loading). The thread count is set here (all cores this process may run on) and the
    count OpenMP then reports is what `cores` says — `torch.distributed.run` exports OMP_NUM_THREADS=1 for
    N > 1, which must not shrink the baseline. If the first step shows that steps + warmup would not fit the
    time budget, the twin is halved per dimension ONCE and `info` (-> the line's `config`) says so.
    Parses the deck through the host-only libjb_host.so: the CUDA library is not mapped by this arm.
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
    try:
        want = len(os.sched_getaffinity(0))
    except AttributeError:
        want = os.cpu_count() or 1
    api.set_num_threads(want)
    cores = int(api.max_threads())

    def make(cells_, particles_):
        pin = ParameterInput(text=workload_text(cells_, min(nb, cells_), particles_))
        # the reference's generator: one XorShift64 pool state per host thread (Kokkos OpenMP backend)
        mb = Mcblock(pin, api=api, cfg_overrides={"rng_mode": capi.JB_RNG_XORSHIFT_THREADS})
        return mb

    def one_step(mb):
        mb.jb.clear_particles()
        mb.ProblemGenerator(density)           # fields + thermal sourcing + tally
        return mb.jb.RadiationStep(0.0, mb.cfg.dt)

    requested = (cells, particles)
    mb = make(cells, particles)
    t0 = time.perf_counter()
    st = one_step(mb)
    t_first = time.perf_counter() - t0
    done_warm = 1
    shrunk = False
    if t_first * (steps + warmup) > budget_s and cells > 32:
        cells, particles = cells // 2, particles // 8
        mb = make(cells, particles)
        done_warm = 0
        shrunk = True
    for _ in range(max(warmup - done_warm, 0)):
        one_step(mb)
    seg = 0
    t0 = time.perf_counter()
    for _ in range(steps):
        st = one_step(mb)
        seg += st["n_segments"]
    el = time.perf_counter() - t0
    info = {"kind": kind, "cores": cores, "flavour": api.flavour().decode(), "cells": cells, "meshblock": min(nb, cells),
            "photons_per_step": particles, "shrunk_to_fit_time_budget": shrunk,
            "requested_twin": {"cells": requested[0], "photons_per_step": requested[1]},
            "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS"),
            "cuda_library_mapped": any("libjaybenne_b200" in l for l in open("/proc/self/maps")),
            "sample": f"{cells}^3-cell twin of the workload (same per-cell loading), {particles} photons, "
                      f"{steps} step(s): thermal sourcing + one cycle, OpenMP over the photons of all blocks on {cores} "
                      f"threads, one XorShift64 pool state per thread as in the reference"}
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
    ap.add_argument("--no-variants", dest="variants", action="store_false",
                    help="skip the `variants` legs (continuing run, sorted continuing run, DDMC at 128^3, free streaming)")
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
        # `config` names what THIS arm ran: a bounded twin of the GPU arm's workload (same per-cell loading,
        # same physics, fewer cells) — the rates are comparable, the problem sizes are not the same
        ref_config = {"workload": workload.replace(f"{cells}^3 cells", f"{info['cells']}^3 cells (CPU twin of the {cells}^3 workload)")
                                          .replace(f"{particles} thermal", f"{info['photons_per_step']} thermal"),
                      "cells": info["cells"], "meshblock": info["meshblock"], "photons_per_step": info["photons_per_step"],
                      "twin_of": {"cells": cells, "photons_per_step": particles},
                      "shrunk_to_fit_time_budget": info["shrunk_to_fit_time_budget"], "requested_twin": info["requested_twin"],
                      "parallelism": f"OpenMP, {info['cores']} threads (omp_get_max_threads), one process",
                      "omp_num_threads_env": info["omp_num_threads_env"], "cuda_library_mapped": info["cuda_library_mapped"],
                      "gray_absorption": "kirchhoff (reference arithmetic)", "rng": "XorShift64 pool, one state per thread"}
        line = {"impl": "reference", "metric": "IMC particle-segments/sec", "value": v, "unit": "segments/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": ref_config,
                "cpu_baseline": {"value": v, "unit": "segments/s", "cores": info["cores"], "kind": info["kind"],
                                 "sample": info["sample"]},
                "e2e": {"value": v, "unit": "segments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    # ------------------------------------------------------------------------- our arm
    import ctypes as C
    import numpy as np
    import torch
    import torch.distributed as dist
    from jaybenne_b200 import Mcblock, ParameterInput, capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: jaybenne_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(vals, op="sum"):
        """all-reduce a list of numbers (float64: exact for counts below 2^53)"""
        if world == 1:
            return [float(v) for v in vals]
        t = torch.tensor([float(v) for v in vals], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return [float(v) for v in t.cpu()]

    class Workload:
        """One configuration resident on this rank's GPU: context, material inputs on the device and in
        pinned host memory, and the step (= one pass of the hot path over one batch)."""

        def __init__(self, cells_, nb_, particles_, stream=False, ddmc=False, sort_interval=0, gray_direct=False):
            self.cells, self.nb, self.particles = cells_, nb_, particles_
            pin = ParameterInput(text=workload_text(cells_, nb_, particles_, stream, ddmc))
            nblocks_total = (cells_ // nb_) ** 3
            over = {"source_nblocks": nblocks_total, "sort_interval": sort_interval,
                    "gray_mode": capi.JB_GRAY_DIRECT if gray_direct else capi.JB_GRAY_KIRCHHOFF}
            # thermal photons + one cycle's emission (particles / nblocks_total, sourcing.cpp:68-69) + slack for
            # the uneven spread of photons over ranks
            cap = int(particles_ * (1.0 + 1.0 / nblocks_total) / world * (1.02 if world == 1 else 1.25)) + 200000
            self.mb = Mcblock(pin, rank=rank, n_ranks=world, capacity=cap, cfg_overrides=over)
            self.jb = jb = self.mb.jb
            if world > 1:
                idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
                if rank == 0:
                    raw = C.create_string_buffer(128)
                    assert jb.api.comm_unique_id(raw) == 0, "ncclGetUniqueId failed"
                    idbuf.copy_(torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8))
                dist.broadcast(idbuf, 0)
                jb.comm_init(bytes(idbuf.cpu().numpy().tobytes()))
            self.ncell = nb_ ** 3
            self.gids = gids = jb.local_gids
            shape = (len(gids), self.ncell)
            self.host_rho = torch.empty(shape, dtype=torch.float64, pin_memory=True)
            self.host_sie = torch.empty(shape, dtype=torch.float64, pin_memory=True)
            self.host_tally = torch.empty(shape, dtype=torch.float64, pin_memory=True)
            self.host_u = torch.empty(shape, dtype=torch.float64, pin_memory=True)
            cv = self.mb.cfg.eos_cv
            for q, gid in enumerate(gids):
                xc = self.mb.mesh.cell_centers(gid)
                rho = density(xc[0][None, None, :], xc[1][None, :, None], xc[2][:, None, None])
                self.host_rho[q] = torch.from_numpy(np.ascontiguousarray(rho).ravel())
                self.host_sie[q] = cv * self.mb.tt0
            self.dev_rho, self.dev_sie = self.host_rho.cuda(), self.host_sie.cuda()
            self.dt = self.mb.cfg.dt
            # per-block addresses are constants of the workload: built once, not inside the timed steps
            self.p = {k: self.ptrs(t) for k, t in (("dev_rho", self.dev_rho), ("dev_sie", self.dev_sie),
                                                   ("host_rho", self.host_rho), ("host_sie", self.host_sie),
                                                   ("host_tally", self.host_tally), ("host_u", self.host_u))}

        def ptrs(self, t):
            return [t[q].data_ptr() for q in range(len(self.gids))]

        def load(self, host):
            jb = self.jb
            jb.clear_particles()
            jb.reset_streams()      # every step sees the same photons: the steps are identical batches
            jb.set_fields_ptr("density", self.gids, self.p["host_rho" if host else "dev_rho"])
            jb.set_fields_ptr("sie", self.gids, self.p["host_sie" if host else "dev_sie"])
            jb.InitializeRadiation(True)
            jb.PostInitialization()

        def step(self, host=False):
            self.load(host)
            st = self.jb.RadiationStep(0.0, self.dt)
            if host:
                self.read_back()
            return st

        def read_back(self):
            self.jb.get_fields_ptr("energy_tally", self.gids, self.p["host_tally"])
            self.jb.get_fields_ptr("internal_energy", self.gids, self.p["host_u"])

        def timed(self, nsteps, host=False):
            """(segments of this rank, device ms, per-step stats)"""
            barrier()
            self.jb.timer_start()
            t0 = time.perf_counter()
            sts = [self.step(host) for _ in range(nsteps)]
            ms = self.jb.timer_stop()
            torch.cuda.synchronize()
            wall = time.perf_counter() - t0
            barrier()
            return sum(st["n_segments"] for st in sts), ms, wall, sts

        def invariants(self, st):
            """Quantities that must not depend on the number of ranks (streams are keyed per photon):
            counts bit for bit, energy sums to rounding (the order of the partial sums differs)."""
            self.read_back()
            dv = (1.0 / self.cells) ** 3
            keys = ("n_start", "n_sourced", "n_absorbed", "n_escaped", "n_end", "n_segments", "n_scatters",
                    "n_block_crossings")
            ekeys = ("e_start", "e_sourced", "e_absorbed", "e_end")
            vals = reduce([st[k] for k in keys] + [st[k] for k in ekeys] +
                          [float(self.host_tally.sum().item()) * dv, float(self.host_u.sum().item())])
            out = {k: int(v) for k, v in zip(keys, vals)}
            for k, v in zip(ekeys + ("tally_sum_times_dv", "internal_energy_sum"), vals[len(keys):]):
                out[k] = v
                out[k + "_hex"] = float(v).hex()
            out["conservation_residual"] = abs(out["e_end"] - out["e_start"] - out["e_sourced"] + out["e_absorbed"]) / max(
                abs(out["e_end"]), 1e-300)
            return out

        def close(self):
            self.jb.close()
            del self.dev_rho, self.dev_sie, self.host_rho, self.host_sie, self.host_tally, self.host_u
            torch.cuda.empty_cache()

    W = Workload(cells, nb, particles, args.stream, args.ddmc, args.sort_interval, args.gray_direct)
    jb, gids, ncell = W.jb, W.gids, W.ncell
    for _ in range(args.warmup):
        W.step()

    # ---- timed: device-resident inputs
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = jb.kernel_launches()
    seg, ms, _, sts = W.timed(args.steps)
    launches = jb.kernel_launches() - launches0
    clocks = sampler.finish()
    tr_ms = sum(st["ms_transport"] for st in sts)
    tr_bytes = cyc_bytes = 0.0
    for st in sts:
        n_in = st["n_start"] + st["n_sourced"] + st["n_received"]
        n_out = st["n_end"] + st["n_sent"]
        tr_bytes += 84.0 * (n_in + n_out) + 32.0 * len(gids) * ncell
        cyc_bytes += 84.0 * (n_in + n_out) + 84.0 * (st["n_start"] + st["n_sourced"]) + 40.0 * len(gids) * ncell
    last = sts[-1]
    ms_max, tr_ms_max = reduce([ms, tr_ms], "max")
    seg_all, = reduce([seg])
    value = seg_all / (ms_max * 1e-3)
    steps_identical = all(st["n_segments"] == sts[0]["n_segments"] and st["n_end"] == sts[0]["n_end"] for st in sts)
    inv = W.invariants(last)
    # device-timed phases of the last step (jb_step_timeline): where a rank's step goes, and how far the ranks differ
    tl = jb.step_timeline()
    phases = {}
    for name, tms in tl:
        phases.setdefault(name, []).append(round(tms, 3))
    tl_sum = {k: sum(v) for k, v in phases.items()}
    tl_keys = ("fields", "source", "transport", "exchange", "stats", "squeeze", "tally", "end")
    tl_max = reduce([tl_sum.get(k, 0.0) for k in tl_keys], "max")
    tl_min = [-v for v in reduce([-tl_sum.get(k, 0.0) for k in tl_keys], "max")]
    timeline = {"rank0_ms": phases, "sum_ms_max_over_ranks": dict(zip(tl_keys, (round(v, 3) for v in tl_max))),
                "sum_ms_min_over_ranks": dict(zip(tl_keys, (round(v, 3) for v in tl_min))),
                "exchange_mode": {0: "none (one rank)", 1: "nccl send/recv per iteration", 2: "direct peer stores over NVLink"}.get(
                    jb.exchange_mode(), "?"),
                "note": "last timed step; a multi-rank step has one transport + exchange pair per iteration, and the exchange "
                        "phase includes waiting for the slowest peer"}

    # ---- timed: end to end through the C ABI with host buffers
    e2e = None
    if not args.no_e2e:
        W.step(host=True)  # warm the host path once
        seg2, _, wall, _ = W.timed(args.steps, host=True)
        el_max, = reduce([wall], "max")
        seg2_all, nb_local = reduce([seg2, len(gids)])
        e2e = {"value": seg2_all / el_max, "unit": "segments/s",
               "h2d_bytes_per_step": int(2 * 8 * ncell * nb_local), "d2h_bytes_per_step": int(2 * 8 * ncell * nb_local),
               "result_checksum": inv["tally_sum_times_dv"],
               "result_checksum_note": "sum over ALL ranks of energy_tally x cell volume (= census energy e_end)"}

    # ---- variants, timed in the same run (each a few seconds; none is the BASELINE metric's config) ----
    def continuing_run(Wk, ncycles):
        """one thermal initialisation followed by consecutive cycles of the same simulation (emission + feedback;
        the population decays by ~1/3 per cycle, holes accumulate and are squeezed by occupancy)"""
        Wk.load(False)
        barrier()
        Wk.jb.timer_start()
        cyc, t_now = [], 0.0
        for _ in range(ncycles):
            st = Wk.jb.RadiationStep(t_now, Wk.dt)
            Wk.jb.UpdateDerivedMaterial()
            t_now += Wk.dt
            cyc.append({"n_start": st["n_start"], "n_segments": st["n_segments"], "ms_total": st["ms_total"],
                        "ms_transport": st["ms_transport"]})
        cms, = reduce([Wk.jb.timer_stop()], "max")
        cseg, = reduce([sum(c["n_segments"] for c in cyc)])
        return {"cycles": ncycles, "value": cseg / (cms * 1e-3), "unit": "segments/s", "ms": cms, "per_cycle_rank0": cyc}

    def variant(name, fn):
        try:
            return fn()
        except Exception as exc:  # noqa: BLE001
            return {"error": f"{type(exc).__name__}: {exc}"}

    variants = {}
    plain = not (args.stream or args.ddmc)
    if args.variants and plain:
        variants["continuing"] = variant("continuing", lambda: dict(
            continuing_run(W, 8), workload=f"{cells}^3: thermal initialisation + 8 consecutive cycles; the pool stays ordered "
                                             "(chunk-local write-back, order-preserving squeeze), no sort"))
        if args.sort_interval == 0:
            def sorted_run():
                Ws = W
                Ws.jb.set_sort_interval(1)
                try:
                    return dict(continuing_run(Ws, 8),
                                workload=f"{cells}^3: same, pool sorted by (meshblock, cell) after every cycle")
                finally:
                    Ws.jb.set_sort_interval(0)
            variants["continuing_sorted"] = variant("continuing_sorted", sorted_run)
    continuing = None
    if args.continuing > 0:
        continuing = continuing_run(W, args.continuing)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    fpeak_err, fpeak = None, None
    try:
        fpeak = jb.measure_fp64_peak()
    except Exception as exc:  # noqa: BLE001
        fpeak_err = str(exc)
    W.close()

    if args.variants and plain:
        def small(stream, ddmc, cells_v, label):
            def run():
                pv = int(round(1.0e9 * (cells_v / 256.0) ** 3))
                Wv = Workload(cells_v, min(nb, cells_v), pv, stream=stream, ddmc=ddmc)
                try:
                    Wv.step()
                    sg, vms, _, vsts = Wv.timed(2)
                    vms_max, vtr = reduce([vms, sum(x["ms_transport"] for x in vsts)], "max")
                    sg_all, = reduce([sg])
                    l = vsts[-1]
                    return {"workload": label.format(cells=cells_v, photons=pv), "value": sg_all / (vms_max * 1e-3),
                            "unit": "segments/s", "ms_per_step": vms_max / 2, "transport_ms_per_step": vtr / 2,
                            "segments_per_photon": l["n_segments"] / max(l["n_start"] + l["n_sourced"], 1),
                            "n_ddmc_segments": l["n_ddmc_segments"]}
                finally:
                    Wv.close()
            return run
        variants["ddmc"] = variant("ddmc", small(False, True, min(cells, 128),
                                   "C5-ddmc {cells}^3 / {photons} photons: sigma*dx = 8, use_ddmc, tau_ddmc = 5 (k_transport_hpool)"))
        variants["stream"] = variant("stream", small(True, False, cells,
                                     "C5-stream {cells}^3 / {photons} photons: dt = dx/(2c), kappa = sigma = 0 (bandwidth-bound limit)"))

    achieved = tr_bytes / (tr_ms * 1e-3) / 1e9 if tr_ms > 0 else 0.0
    # DRAM traffic of the kernel: measured once with ncu on the 64^3 twin (profiles/*_traffic.json) as a ratio to
    # the algorithmic bytes of that launch; reported here scaled to this run's launch (None if the file is absent)
    traffic, traffic_src, ncu_pipe, tj = None, None, {}, {}
    for tname in ("r02_traffic.json", "r01_traffic.json"):
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", tname)))
            break
        except Exception:
            continue
    if tj and plain:
        traffic = tj["ratio"] * tr_bytes / max(args.steps, 1)
        traffic_src = f"ncu at {tj['cells']}^3 ({tname}): dram bytes / algorithmic bytes = {tj['ratio']}, scaled to this launch"
        ncu_pipe = {"fp64_pipe_pct": tj.get("fp64_pipe_pct"), "issue_active_pct": tj.get("issue_active_pct"),
                    "kernel": tj.get("kernel")}
    fps = float(tj.get("fp64_flops_per_segment", 324.0)) if tj else 324.0
    spp = seg / max(args.steps * (last["n_start"] + last["n_sourced"]), 1)
    kernel_name = (os.environ.get("JB_TRANSPORT_VARIANT", "2") == "2" and "k_transport_hpool" or "k_transport_hybrid") if args.ddmc else os.environ.get("JB_TRANSPORT_VARIANT", "2") == "1" and "k_transport_imc" or "k_transport_pool"
    roofline = {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "ncu": ncu_pipe,
                "note": f"{spp:.1f} segments per photon make this kernel FP64-issue / latency bound ({fps:.0f} FP64 flop per "
                        f"segment against {tr_bytes / max(seg, 1):.0f} algorithmic bytes); variants.stream measures the same "
                        "kernel in the bandwidth-bound regime",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                "bytes_per_launch": tr_bytes / max(args.steps, 1), "ms_per_launch": tr_ms / max(args.steps, 1),
                "bytes_per_launch_formula": "84 B x (photons in + photons out) + 32 B x cells (the CellProps record the "
                                            "kernel reads; SURVEY 8d's 40 B x cells is per CYCLE and is used in whole_cycle_GBps)",
                "bytes_per_segment": tr_bytes / max(seg, 1),
                "whole_cycle_GBps": cyc_bytes / (ms * 1e-3) / 1e9,
                "segments_per_photon": spp}

    # second roofline (SURVEY 8d honesty note): FP64 pipe. Peak = DFMA chain kernel measured now on this GPU;
    # flops per segment = DFMA x2 + DMUL + DADD of the kernel's main body, counted by ncu (profiles/*_traffic.json)
    if fpeak:
        fach = (seg / (tr_ms * 1e-3)) * fps / 1e12 if tr_ms > 0 else 0.0
        roofline["fp64"] = {"bound": "fp64", "peak": fpeak, "unit": "TFLOP/s", "peak_source": "k_fp64_chain measured in this run",
                            "flops_per_segment": fps, "achieved": fach, "frac": fach / fpeak if fpeak > 0 else None,
                            "pipe_active_pct_ncu": ncu_pipe.get("fp64_pipe_pct")}
    else:
        roofline["fp64"] = {"error": fpeak_err}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_cells = min(args.cpu_cells, cells)
        cpu_particles = int(round(particles * (cpu_cells / cells) ** 3))
        v, cms, info = cpu_arm(cpu_cells, nb, cpu_particles, 1, 1)
        cpu = {"value": v, "unit": "segments/s", "cores": info["cores"], "kind": info["kind"], "sample": info["sample"],
               "ms_per_step": cms, "twin": {k: info[k] for k in ("cells", "photons_per_step", "shrunk_to_fit_time_budget")}}

    if rank == 0:
        line = {"metric": "IMC particle-segments/sec", "value": value, "unit": "segments/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / max(args.steps, 1),
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config, "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
                "roofline": roofline, "cpu_baseline": cpu,
                "transport_ms_per_step": tr_ms_max / max(args.steps, 1),
                "invariants": dict(inv, steps_identical=steps_identical,
                                   note="all-reduced over the ranks; every step sources the same photons (jb_reset_streams), "
                                        "so these are equal for every N, K and W"),
                "timeline": timeline, "variants": variants or None, "continuing": continuing,
                "last_step_rank0": {k: last[k] for k in ("n_start", "n_sourced", "n_absorbed", "n_end", "n_sent", "n_received",
                                                          "n_segments", "n_scatters", "n_block_crossings", "n_iterations")}}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
