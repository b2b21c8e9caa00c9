"""Print the interesting parts of a bench.py JSON line."""
import json
import signal
import sys

signal.signal(signal.SIGPIPE, signal.SIG_DFL)

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
if "value" in d:
    print("value", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms/step", round(d["ms_per_step"], 3),
          "serial", round(d.get("roofline", {}).get("serial_ms_per_step", 0), 3), "cpu", (d.get("cpu_baseline") or {}).get("value"))
if "stages" in d:
    print({k: round(v["ms_per_step"], 3) for k, v in d["stages"].items()})
    r = d["roofline"]
    print("roofline", r["kernel"], round(r["frac"], 3), "xu", round(r["xu_pipe"]["frac"], 3), d["clocks"], d["run"]["ransac_iterations"])
for k in ("stream", "cfg4", "hard_ransac", "cfg2"):
    if k in d:
        v = d[k]
        if k == "cfg4":
            print("cfg4 value", round(v["value"], 1), "e2e", round(v["e2e"]["value"], 1), {s: round(x["ms_per_step"], 3) for s, x in v["stages"].items()}, "cpu", (v.get("cpu_baseline") or {}).get("value"))
        else:
            print(k, json.dumps(v)[:1500])
