"""Reads `ncu --page source --csv --print-source sass` output and prints the top-N SASS lines by
stall samples with their dominant stall reason (developer aid for reading .ncu-rep files)."""
import csv, sys
path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30
rows = list(csv.reader(open(path)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
tot = 0
for n, r in enumerate(rows[2:]):
    if len(r) < len(hdr):
        continue
    s = int(r[ix["# Samples"]] or 0)
    tot += s
    data.append((s, n, r))
print("total samples", tot)
for s, n, r in sorted(data, reverse=True)[:top]:
    reasons = sorted(((int(r[ix[c]] or 0), c) for c in stall_cols), reverse=True)[:2]
    print(f"{s:7d} {100*s/tot:5.1f}% line {n:5d} {r[ix['Source']].strip()[:70]:70s} exec={r[ix['Instructions Executed']]:>9s} "
          + " ".join(f"{c[6:]}={v}" for v, c in reasons if v))
