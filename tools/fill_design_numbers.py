"""Writes the round's bench numbers (profiles/r02_bench_{1,8}gpu.json) into DESIGN.md's table and README."""
import json, os, re
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
d1 = json.loads(open(os.path.join(ROOT, "profiles/r02_bench_1gpu.json")).read())
d8 = json.loads(open(os.path.join(ROOT, "profiles/r02_bench_8gpu.json")).read())
def k(x): return "%.1f k" % (x / 1000.0) if x >= 10000 else "%d" % round(x)
def cfg(d, c):
    v = d["configs"][c]
    if c == "c3":
        r = v["regimes"]
        return "%s (%.2f of HBM) / %s queries/s" % (k(r[0]["value"]), r[0]["hbm_frac"], k(r[1]["value"]))
    if c == "c4":
        return "%s / %s queries/s" % (k(v["value"]), k(v["e2e"]["value"]))
    if c == "c5":
        return k(v["value"])
    return "%s queries/s (%.2f of POPC issue)" % (k(v["value"]), v["regimes"][0].get("popc_issue_frac", 0))
rows = [
 ("headline, device-resident", "%s queries/s, %.2f of HBM" % (k(d1["value"]), d1["roofline"]["frac"]), "%s queries/s, %.2f of HBM on the 12.5 M-row shards" % (k(d8["value"]), d8["roofline"]["frac"])),
 ("headline, end to end", k(d1["e2e"]["value"]), k(d8["e2e"]["value"])),
 ("`throughput_mode` (library default: %d / %d queries per pass)" % (round(d1["throughput_mode"]["queries_per_pass"]), round(d8["throughput_mode"]["queries_per_pass"])), k(d1["throughput_mode"]["value"]), k(d8["throughput_mode"]["value"])),
 ("c2", cfg(d1, "c2"), cfg(d8, "c2")),
 ("c3 at 2 per pass / 24 per pass", cfg(d1, "c3"), cfg(d8, "c3")),
 ("c4 (device-resident / end to end)", cfg(d1, "c4"), cfg(d8, "c4")),
 ("c5", cfg(d1, "c5") + " `add`/s", cfg(d8, "c5") + " `add`/s over 8 replicas, MIX included"),
 ("CPU engine, same table (%d cores)" % d1["cpu_baseline"]["cores"], "%.1f queries/s" % d1["cpu_baseline"]["value"], "—"),
]
table = "| | 1 GPU | 8 GPUs |\n|---|---|---|\n" + "".join("| %s | %s | %s |\n" % r for r in rows)
p = os.path.join(ROOT, "DESIGN.md")
s = open(p).read()
a = s.index("| | 1 GPU | 8 GPUs |")
b = s.index("\n\n", a)
s = s[:a] + table.rstrip("\n") + s[b:]
open(p, "w").write(s)
print(table)
