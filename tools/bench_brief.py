"""one-line digest of a bench.py log: python tools/bench_brief.py <log>"""
import json, sys
try:
    j = json.loads([x for x in open(sys.argv[1]) if x.startswith("{")][-1])
    k = j["roofline"].get("kernels", {})
    print("ms=%.4f" % j["ms_per_step"], "e2e_ms=%.4f" % (j["config"]["global_batch"] / j["e2e"]["value"] * 1e3),
          " ".join("%s=%.1fus" % (n.split("_kernel")[0], v["ms_per_step"] * 1e3) for n, v in k.items()),
          "loss=%.7f" % j["loss"])
except Exception as e:  # noqa: BLE001
    print("no bench line:", repr(e)[:80])
