"""Reads bench.py's JSON line from stdin and prints the fields the sweeps compare."""
import json
import sys

for line in sys.stdin:
    if line.startswith("{"):
        d = json.loads(line)
        an = d.get("step_anatomy_us") or {}
        print("us/step %.2f | M/s %.3f | gather: events %.2f us (frac %.3f), in-step %s us | update %s | sample %s | period %s | eager M/s %.3f"
              % (d["ms_per_step"] * 1e3, d["value"] / 1e6, d["roofline"]["avg_launch_us"], d["roofline"]["frac"], an.get("gather_grid"),
                 an.get("update"), an.get("sample"), an.get("period"), (d.get("value_eager_launches") or 0) / 1e6))
