"""Per kernel of an `ncu -i X.ncu-rep --page source --csv` dump: stall-reason totals and the SASS lines with most samples.
usage: python tools/ncu_top_stalls.py dump.csv [top_n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = [r]
        blocks.append(cur)
    elif cur is not None:
        cur.append(r)
for blk in blocks:
    print(blk[0][1])
    hdr = blk[1]
    iS, iSrc, iEx = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
    stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    body = [r for r in blk[2:] if len(r) == len(hdr)]
    tot = sum(int(r[iS] or 0) for r in body) or 1
    agg = {}
    for r in body:
        for i, h in stall_cols:
            agg[h] = agg.get(h, 0) + int(r[i] or 0)
    print("  total samples", tot, " ".join("%s=%.1f%%" % (h[6:], 100.0 * v / tot) for v, h in sorted(((v, h) for h, v in agg.items()), reverse=True)[:8]))
    for r in sorted(body, key=lambda r: -int(r[iS] or 0))[:top_n]:
        st = sorted(((int(r[i] or 0), h) for i, h in stall_cols), reverse=True)[:2]
        print("  %5.1f%% ex=%-9s %-64s %s" % (100.0 * int(r[iS]) / tot, r[iEx], r[iSrc].strip()[:64], ", ".join("%s=%d" % (h[6:], v) for v, h in st if v)))
