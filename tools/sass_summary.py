"""Per-kernel SASS instruction counts of the Blackwell-native paths in libtt_b200.so (cuobjdump -sass; runs without a GPU).
python tools/sass_summary.py > profiles/sass_summary.txt"""
import collections, hashlib, re, subprocess, sys
so = sys.argv[1] if len(sys.argv) > 1 else "two_towers_b200/libtt_b200.so"
cols = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "UBLKPF", "UTCBAR", "SYNCS", "HMMA", "MUFU"]
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
names = {}
try:
    dem = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.split("\n")
    for m, d in zip(re.findall(r"Function : (\S+)", out), dem):
        names[m] = d
except OSError:
    pass
kern, cnt, tot = None, collections.defaultdict(collections.Counter), collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1).split(".")[0]
        tot[kern] += 1
        if op in cols:
            cnt[kern][op] += 1
sha = hashlib.sha256(open(so, "rb").read()).hexdigest()[:16]
print(f"cuobjdump -sass {so} (sha256 {sha}..., sm_100a only): per-kernel SASS instruction counts of the Blackwell-native paths")
print("UTCHMMA = tcgen05.mma kind::f16 (cta_group::1 and ::2), LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA tensor load/store, UTMAPF/UBLKPF = L2 prefetch,")
print("UBLKCP = cp.async.bulk (DSMEM copy), UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, HMMA = legacy mma.sync (none expected), MUFU = ex2/rcp.")
print("Kernels without any of the first nine are omitted.\n")
print("  total " + " ".join(f"{c:>8s}" for c in cols) + "  kernel")
for k in sorted(tot, key=lambda k: -tot[k]):
    if not any(cnt[k][c] for c in cols[:9]):
        continue
    print(f"{tot[k]:7d} " + " ".join(f"{cnt[k][c]:8d}" for c in cols) + "  " + names.get(k, k))
