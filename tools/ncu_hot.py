"""Top stall sites of one kernel in an ncu report (SASS level, with samples). Usage: ncu_hot.py report.ncu-rep [kernel-substr] [topN]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
sub = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(raw)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": [], "hdr": None}
        blocks.append(cur)
    elif cur is not None and row and row[0] == "Address":
        cur["hdr"] = row
    elif cur is not None and cur["hdr"] and len(row) == len(cur["hdr"]):
        cur["rows"].append(row)
for b in blocks:
    if sub not in b["name"] or b["name"] in globals().get("_seen", set()):
        continue
    h = b["hdr"]
    isamp, isrc, iex = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
    tot = sum(int(r[isamp]) for r in b["rows"]) or 1
    print("==", b["name"][:100], "samples", tot, "instructions", len(b["rows"]))
    # cumulative by position to see phases
    acc = 0
    marks = []
    for k, r in enumerate(b["rows"]):
        acc += int(r[isamp])
        if any(t in r[isrc] for t in ("BAR.SYNC", "EXIT")):
            marks.append((k, r[isrc].strip()[:30], acc))
    prev = 0
    for k, s, a in marks:
        print(f"  up to instr {k:5d} {s:32s} {100*(a-prev)/tot:5.1f}% of samples")
        prev = a
    order = sorted(range(len(b["rows"])), key=lambda k: -int(b["rows"][k][isamp]))[:top]
    for k in sorted(order):
        r = b["rows"][k]
        print(f"  {k:5d} {100*int(r[isamp])/tot:5.1f}%  exec {r[iex]:>8s}  {r[isrc].strip()[:90]}")
    seen = globals().setdefault("_seen", set())
    if sub or len(seen) > 12:
        break
    seen.add(b["name"])
