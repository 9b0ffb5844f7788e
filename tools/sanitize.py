#!/usr/bin/env python
"""compute-sanitizer memcheck + racecheck over small decks run by mcblock_b200 (the reference's only
guard is PARTHENON_DEBUG_REQUIRE, transport.cpp:100-111,152-153; the kernels here use shared-memory pools,
shared cursors and outbox atomics). Usage: python tools/sanitize.py <outdir> [deck ...]
Writes <outdir>/sanitizer_<tool>_<deck>.log and a summary line per run on stdout."""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from jaybenne_b200.decks import CI_OVERRIDES, deck_text  # noqa: E402

BIN = os.path.join(ROOT, "jaybenne_b200", "mcblock_b200")
# a few cycles are enough: every kernel of the cycle runs in each one
EXTRA = {"smr3d_hybrid": ["--max-cycles", "2"]}
DEFAULT_EXTRA = ["--max-cycles", "3"]


def main():
    out = sys.argv[1]
    decks = sys.argv[2:] or ["inf", "stepdiff_smr_hybrid", "smr3d_hybrid"]
    os.makedirs(out, exist_ok=True)
    rc_all = 0
    with tempfile.TemporaryDirectory() as tmp:
        for name in decks:
            deck = os.path.join(tmp, name + ".in")
            with open(deck, "w") as f:
                f.write(deck_text(name))
            args = [BIN, "-i", deck, "-d", os.path.join(tmp, name + "_out")]
            args += [f"{b}/{k}={v}" for b, k, v in CI_OVERRIDES.get(name, [])] + EXTRA.get(name, DEFAULT_EXTRA)
            for tool in ("memcheck", "racecheck"):
                log = os.path.join(out, f"sanitizer_{tool}_{name}.log")
                cmd = ["compute-sanitizer", "--tool", tool, "--error-exitcode", "9", "--print-limit", "20"] + args
                r = subprocess.run(cmd, capture_output=True, text=True, timeout=1500)
                with open(log, "w") as f:
                    f.write("command: " + " ".join(cmd[:6]) + " mcblock_b200 -i %s.in " % name + " ".join(args[5:]) + "\n")
                    f.write(r.stdout[-6000:] + "\n--- stderr ---\n" + r.stderr[-6000:])
                tail = [l for l in (r.stdout + r.stderr).splitlines() if "SUMMARY" in l]
                print(name, tool, "rc", r.returncode, tail[-1] if tail else "(no summary line)")
                rc_all |= r.returncode
    sys.exit(1 if rc_all else 0)


if __name__ == "__main__":
    main()
