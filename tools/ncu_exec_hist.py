"""Executed-instruction budget of a capture by how often each SASS line ran (run where the report is):
python tools/ncu_exec_hist.py rep"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]; data = rows[2:]
ci = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ci["Instructions Executed"]]) for r in data)
b = collections.Counter(); n = collections.Counter()
for r in data:
    e = int(r[ci["Instructions Executed"]])
    b[e] += e; n[e] += 1
print("total executed", tot)
for e, s in sorted(b.items(), key=lambda x: -x[1])[:14]:
    print("%12d executions x %4d SASS lines = %5.1f %% of all" % (e, n[e], 100.0 * s / tot))
spin = [r for r in data if "TRYWAIT" in r[ci["Source"]]]
for r in spin[:12]:
    print("trywait", r[ci["Instructions Executed"]], r[ci["# Samples"]])
