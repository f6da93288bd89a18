#!/usr/bin/env python3
"""Short view of a bench.py JSON line (the last JSON line of the file)."""
import json
import signal
import sys

signal.signal(signal.SIGPIPE, signal.SIG_DFL)

for path in sys.argv[1:]:
    line = [l for l in open(path).read().splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    print(path, "N=%s" % d.get("n_gpus"), "value %.4g" % d["value"], "ms %.3f" % d["ms_per_step"],
          "e2e %.4g" % d["e2e"]["value"], "frac %.3f" % d["roofline"]["frac"])
    print("  parity", {k: v for k, v in d["parity"].items() if k != "what"})
    for k, c in (d.get("configs") or {}).items():
        if c is None or "rows" in c:
            continue
        b2b = (c.get("back_to_back") or {}).get("ms_per_step")
        print("  %-16s ms %.3f  b2b %s  value %.4g  frac %.3f" % (
            k, c["ms_per_step"], "%.3f" % b2b if b2b else "-", c["value"], c["roofline"]["frac"] or 0))
        if k.startswith("bh"):
            st = c["roofline"]["stages"]
            print("     stages", {a: round(b["ms"], 3) for a, b in st.items()},
                  "trav_min", round(st["traversal"].get("ms_min_over_ranks", 0), 3))
            print("     e2e %.4g" % c["e2e"]["value"], "parity",
                  {a: b for a, b in c["parity"].items() if a != "what"})
