"""Summarise `ncu --page source --csv` output: stall mix and the hottest instructions (by warp-stall samples)."""
import csv, sys
path, topn = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30
rows = list(csv.reader(open(path)))
hdr = next(r for r in rows if r and r[0] == 'Address')
idx = {h: i for i, h in enumerate(hdr)}
def num(r, k):
    try: return int(float(r[idx[k]] or 0))
    except ValueError: return 0
seen, data = set(), []
for r in rows:
    if len(r) != len(hdr) or r[0] == 'Address' or r[0] in seen: continue
    seen.add(r[0]); data.append(r)
tot = sum(num(r, '# Samples') for r in data)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {s: sum(num(r, s) for r in data) for s in stalls}
print("samples", tot, "instructions", len(data))
print(", ".join(f"{k[6:]} {v / max(tot, 1):.2f}" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
for r in sorted(data, key=lambda r: -num(r, '# Samples'))[:topn]:
    st = sorted(((s[6:], num(r, s)) for s in stalls), key=lambda x: -x[1])[:2]
    print(str(num(r, '# Samples')).rjust(6), r[idx['Address']][-5:], r[idx['Source']][:84].ljust(84), st)
