"""Print the headline numbers of a bench.py JSON line file: python tools/show_bench.py file.json"""
import json
import sys

for l in open(sys.argv[1]):
    l = l.strip()
    if l.startswith("{"):
        d = json.loads(l)
        print("value %.1f %s  ms/step %.3f  e2e %.1f  launches %s" % (d["value"], d["unit"], d["ms_per_step"],
                                                                       d.get("e2e", {}).get("value", float("nan")), d.get("gpu_launches")))
