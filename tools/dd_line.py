"""stdin: the JSON line of tools/bench_domain.py; prints the few fields an A/B needs.  usage: ... | python tools/dd_line.py <label>"""
import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(sys.argv[1] if len(sys.argv) > 1 else "", "ms/slice", round(d["ms_per_slice"], 4), "value %.4g" % d["value"], "exch", d.get("exchanges"),
      "redund", round(d.get("redundancy", 0), 4), "rebuilds", d.get("rebuilds"), json.dumps(d.get("timeline_ms_per_slice_max_over_ranks")))
