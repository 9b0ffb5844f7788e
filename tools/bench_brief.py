"""one line per bench JSON file: value, ms per step, transport ms, iterations, invariants"""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f).read().strip().splitlines() if l.startswith("{")][-1])
        v = d.get("variants") or {}
        print(f.split("/")[-1], "value %.4g" % d["value"], "ms/step %.1f" % d["ms_per_step"],
              "transport %.1f" % d["transport_ms_per_step"], "iters", d["last_step_rank0"]["n_iterations"],
              "n_seg", d["invariants"]["n_segments"], "n_end", d["invariants"]["n_end"], d["invariants"]["e_end_hex"],
              " ".join("%s %.4g" % (k, x["value"]) for k, x in v.items() if isinstance(x, dict) and "value" in x))
    except Exception as e:  # noqa: BLE001
        print(f, "FAILED", e)
