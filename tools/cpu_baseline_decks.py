#!/usr/bin/env python
"""CPU baseline beside the GPU numbers: the reference's per-particle headers
driven by the oracle (oracle/_ref timing build when present, else the oracle port's -O3 build), OpenMP over photons
with one XorShift64 pool state per thread, timed on 1 host core and on all of them for

  C1  inputs/inf.in as shipped (100 cycles),
  C2  inputs/stepdiff.in with the CI overrides (cycles bounded, see --c2-cycles),
  C5  the 64^3 twin of the synthetic workload (thermal sourcing + one cycle, as bench.py steps it).

Reported, not a target. Each (config, threads) pair runs in its own process because OMP_NUM_THREADS is read when the
library loads.  usage: python tools/cpu_baseline_decks.py [--out profiles/r01_cpu_baseline_decks.json]
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def child(config, cycles):
    import oracle_binding as ob
    import parity_util as pu
    import bench
    from jaybenne_b200 import Mcblock, ParameterInput, capi
    ref_fast = os.path.join(ob.ORACLE_DIR, "_ref", "libjb_oracle_ref_fast.so")
    if os.path.exists(ref_fast):
        ob.FLAVOURS["ref_fast"] = ref_fast
        flavour, kind = "ref_fast", "reference"
    else:
        flavour, kind = "fast", "port"
    api = ob.oracle_api(flavour)
    over = {"rng_mode": capi.JB_RNG_XORSHIFT_THREADS}
    if config == "C5":
        cells, particles = 64, 15625000
        mb = Mcblock(ParameterInput(text=bench.workload_text(cells, 32, particles)), api=api, cfg_overrides=over)
        t0 = time.perf_counter()
        seg = 0
        for _ in range(cycles):
            mb.jb.clear_particles()
            mb.ProblemGenerator(bench.density)
            seg += mb.jb.RadiationStep(0.0, mb.cfg.dt)["n_segments"]
        el = time.perf_counter() - t0
        what = f"64^3 cells, {particles} photons, {cycles} x (thermal sourcing + one cycle)"
    else:
        name = {"C1": "inf", "C2": "stepdiff"}[config]
        mb = Mcblock(pu.make_pin(name, ci=(config == "C2")), api=api, cfg_overrides=over)
        mb.ProblemGenerator()
        t0 = time.perf_counter()
        mb.Execute(max_cycles=cycles)
        el = time.perf_counter() - t0
        seg = sum(h["n_segments"] for h in mb.history)
        cycles = len(mb.history)
        what = f"inputs/{name}.in" + (" with the CI overrides" if config == "C2" else "") + f", {cycles} cycles"
    print(json.dumps({"config": config, "what": what, "kind": kind, "flavour": api.flavour().decode(),
                      "threads": int(os.environ.get("OMP_NUM_THREADS", "0")), "segments": int(seg),
                      "seconds": round(el, 3), "segments_per_s": seg / el, "ms_per_cycle": 1e3 * el / max(cycles, 1)}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--child", nargs=2)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r01_cpu_baseline_decks.json"))
    ap.add_argument("--c2-cycles", type=int, default=2)
    a = ap.parse_args()
    if a.child:
        return child(a.child[0], int(a.child[1]))
    ncpu = os.cpu_count() or 1
    rows = []
    for config, cycles in (("C1", 100), ("C2", a.c2_cycles), ("C5", 1)):
        for threads in sorted({1, ncpu}):
            env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close")
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", config, str(cycles)], env=env,
                               capture_output=True, text=True, check=True)
            rows.append(json.loads(r.stdout.strip().splitlines()[-1]))
            print(rows[-1], flush=True)
    cpu = ""
    try:
        cpu = [l.split(":", 1)[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
    except Exception:
        pass
    with open(a.out, "w") as f:
        json.dump({"host": {"cpu": cpu, "logical_cores": ncpu, "where": "build container (no GPU)"}, "rows": rows}, f,
                  indent=1)
        f.write("\n")


if __name__ == "__main__":
    main()
