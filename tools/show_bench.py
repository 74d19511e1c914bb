"""Development helper: one-line summaries of bench.py JSON lines."""
import json
import sys

for f in sys.argv[1:]:
    t = open(f).read().strip().splitlines()
    if not t:
        print(f, "EMPTY")
        continue
    d = json.loads(t[-1])
    if d.get("op") in ("batch_exp", "batch_to_special"):
        print(f, d["op"], "e2e M/s", round(d["value"] / 1e6, 1), "ms", round(d["ms_per_step"], 2), "kernel_ms",
              round(d["kernel_ms"], 3), "frac", round(d["roofline"]["frac"], 3), "cpu M/s",
              round(d["cpu_baseline"]["value"] / 1e6, 2), d["verified"]["ok"],
              {k: round(v, 2) for k, v in d["phases_ms"].items() if v > 0.01})
        continue
    if d.get("impl") == "reference":
        print(f, "REF M/s", round(d["value"] / 1e6, 3), "ms", round(d["ms_per_step"], 1), d["steps"], d["verified"])
        continue
    print(f, "N", d["n_gpus"], "ms", round(d["ms_per_step"], 3), "Mpts/s", round(d["value"] / 1e6, 1), "c", d["window_bits"],
          {k: round(v, 3) for k, v in d["phases_ms"].items()})
    e = d.get("e2e")
    if e:
        print("    e2e ms", round(e["ms_per_step"], 2), "M/s", round(e["value"] / 1e6, 1), "| resident_bases ms",
              round(e["resident_bases"]["ms_per_step"], 2), "| pageable M/s", round(e["pageable_value"] / 1e6, 1))
    s = d.get("strong")
    print("    verified", d["verified"]["paths"], "| strong", s and (s["points_total"], round(s["ms_per_step"], 3),
          round(s["value"] / 1e6, 1)), "| launches", d["gpu_launches"], "| frac", round(d["roofline"]["frac"], 3),
          "exec", round(d["roofline"]["executed"]["frac"], 3))
