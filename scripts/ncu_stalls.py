"""Top stall sites of a kernel from an ncu report (source page).  python scripts/ncu_stalls.py X.ncu-rep [N]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, data = rows[1], rows[2:]
iS, iSrc, iEx = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[iS] or 0) for r in data)
agg = {}
for r in data:
    for i in stall:
        agg[hdr[i][6:]] = agg.get(hdr[i][6:], 0) + int(r[i] or 0)
print(rows[0][1][:90], " samples", tot, " instr", sum(int(r[iEx] or 0) for r in data))
print("by reason:", [(k, round(100 * v / tot, 1)) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]])
for r in sorted(data, key=lambda r: -int(r[iS] or 0))[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    st = sorted(((hdr[i][6:], int(r[i] or 0)) for i in stall if int(r[i] or 0) > 0), key=lambda x: -x[1])[:2]
    print(f"{int(r[iS]):7d} {100 * int(r[iS]) / tot:5.1f}%  ex={int(r[iEx] or 0):9d}  {r[iSrc][:64]:64s} {st}")
