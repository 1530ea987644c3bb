"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list: ms per step, launches, mean us."""
import csv, re, collections, sys
path, steps = sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else 3.0
rows = list(csv.DictReader(l for l in open(path) if not l.startswith('==')))
agg = collections.defaultdict(list)
for r in rows:
    nm = re.sub(r'\(.*', '', r['Kernel Name']).replace('void ', '')
    if not nm.startswith('tt::'):
        nm = 'torch:' + nm[:44]
    agg[(nm, r['Grid Size'])].append(float(r['Metric Value']) / 1e3)
probe = ('tt::gemm_f32_kernel<1, 1>', '(6, 4096, 1)')
tot = sum(sum(v) for k, v in agg.items() if k != probe)
print(f"total {tot / 1e3 / steps:.2f} ms/step (excluding the stand-alone roofline probe)")
for k, v in sorted(agg.items(), key=lambda x: -sum(x[1]))[:int(sys.argv[3]) if len(sys.argv) > 3 else 24]:
    print(f"{sum(v) / 1e3 / steps:8.2f} ms/step n/step={len(v) / steps:6.1f} mean {sum(v) / len(v):9.1f} us  {k[0]} {k[1]}")
