#!/usr/bin/env python
"""Generates the golden fixtures in this directory FROM THE REFERENCE'S OWN CODE.

Run in the development container only (it needs /root/reference):

    python tests/golden/make_golden.py

It drives oracle/_ref/libjb_oracle_ref.so — the flavour of the oracle whose per-particle
physics is the reference's own headers compiled verbatim where they lie
(src/jaybenne/transport_utils.hpp, scattering.hpp, planck.hpp, src/mcblock/opacity.hpp and
eos.hpp -> singularity-opac / singularity-eos; recipe: oracle/Makefile `ref`) — and records

  closures.json      EOS / opacity / scattering / emissivity values (singularity closures)
  steps.json         ptcl_transport_step(+scatter), ptcl_ddmc_albedo + ptcl_ddmc_step and
                     sample_Planck_energy on seeded random inputs, inputs and outputs as
                     IEEE-754 hex so the comparison is bit for bit
  runs.json          per-cycle counts / energies and a digest of the census swarm for whole
                     runs of the shipped decks (inf, stepdiff, stepdiff_ddmc, the SMR decks)

The files are small and committed; tests/test_oracle_pin.py checks the restated oracle
(the twin of the CUDA path) against them on any machine, without /root/reference.
"""
import ctypes as C
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import golden_cases as gc  # noqa: E402
from oracle_binding import oracle_api  # noqa: E402


def main():
    if not os.path.isdir("/root/reference/src/jaybenne"):
        raise SystemExit("needs /root/reference (development container only)")
    api = oracle_api("ref")
    assert api.flavour() == b"reference-headers+libm", api.flavour()
    out = {
        "closures.json": gc.closure_table(api),
        "steps.json": gc.step_table(api),
        "runs.json": gc.run_table(api),
    }
    for name, data in out.items():
        data["_generated_by"] = "tests/golden/make_golden.py with oracle/_ref/libjb_oracle_ref.so (reference headers + libm)"
        with open(os.path.join(HERE, name), "w") as f:
            json.dump(data, f, indent=1, sort_keys=True)
            f.write("\n")
        print("wrote", name, os.path.getsize(os.path.join(HERE, name)), "bytes")


if __name__ == "__main__":
    main()
