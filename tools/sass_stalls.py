"""Per-instruction stall listing of a range of an `ncu --page source --csv --print-source sass` dump."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
recs = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name": break
    if len(r) < len(hdr): continue
    recs.append(r)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
a, b = int(sys.argv[2]), int(sys.argv[3])
tot = 0
for k in range(a, min(b, len(recs))):
    r = recs[k]
    s = int(r[ix['# Samples']] or 0); tot += s
    st = sorted(((int(r[ix[h]] or 0), h[6:]) for h in stalls), reverse=True)[:3]
    print(f"{k:5d} {s:4d} {r[ix['Source']][:64]:64s} " + " ".join(f"{h}:{n}" for n, h in st if n))
print("total samples", tot, "of", sum(int(r[ix['# Samples']] or 0) for r in recs))
