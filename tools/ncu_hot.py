"""Top stalled SASS instructions of a capture (run here): python tools/ncu_hot.py rep [n]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]; data = rows[2:]
ci = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[ci["# Samples"]]) for r in data)
print("total samples", tot, "instructions executed", sum(int(r[ci["Instructions Executed"]]) for r in data))
agg = {s: sum(int(r[ci[s]]) for r in data) for s in stalls}
print(sorted(agg.items(), key=lambda x: -x[1])[:8])
for r in sorted(data, key=lambda r: -int(r[ci["# Samples"]]))[:n]:
    st = {s: int(r[ci[s]]) for s in stalls if int(r[ci[s]]) > 0}
    print(r[ci["# Samples"]], r[ci["Instructions Executed"]], r[ci["Source"]].strip()[:58], sorted(st.items(), key=lambda x: -x[1])[:3])
