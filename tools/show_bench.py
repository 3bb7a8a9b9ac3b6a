"""Print the headline fields of bench.py JSON lines: show_bench.py file..."""
import json
import sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable:", e)
        continue
    print("%s: n=%d %.3f ms/step %.0f %s | %s | %s" % (
        f, d.get("n_gpus", 1), d["ms_per_step"], d["value"], d["unit"],
        {k: round(v, 3) for k, v in d.get("kernels_ms", {}).items()} or d.get("impl", ""),
        str(d.get("config", {}).get("parallelism", ""))[:48]))
