"""Prints the SASS instructions with the most warp-stall samples from an `ncu --page source --csv` dump."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
iS, iN, iSrc = hdr.index("# Samples"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Source")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) == len(hdr)]
tot = sum(int(r[iS] or 0) for r in body)
print("total samples", tot)
top = sorted(body, key=lambda r: -int(r[iS] or 0))[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for r in top:
    st = sorted(((int(r[i] or 0), h) for i, h in stall_cols), reverse=True)[:2]
    print("%5.1f%%  %-60s %s" % (100.0 * int(r[iS]) / tot, r[iSrc].strip()[:60], ", ".join("%s=%d" % (h[6:], v) for v, h in st if v)))
