#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of libwhisper_b200.so (cuobjdump -sass): the Blackwell-native claim, checkable without rebuilding.
UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st (TMEM), UTMALDG / UTMASTG / UTMAREDG = TMA tensor load / store / reduce,
UBLKCP = cp.async.bulk, SYNCS = mbarrier ops, UTCBAR = tcgen05.commit, UTCATOMSWS = tcgen05.alloc / dealloc."""
import collections, re, subprocess, sys
lib = sys.argv[1] if len(sys.argv) > 1 else "whisper_rust_b200/lib/libwhisper_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
watch = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UTCBAR", "UTCATOMSWS", "SYNCS", "HMMA", "MUFU", "LDG", "STG", "LDS", "STS", "FFMA", "DADD", "DFMA"]
kern, hist = None, collections.OrderedDict()
arch = None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().replace("(anonymous namespace)::", "").split("(")[0]
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s*arch = (\S+)", line)
    if m:
        arch = m.group(1)
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and kern:
        op = m.group(1)
        hist[kern][op.split(".")[0]] += 1
print(f"# {lib}: {arch}; instruction counts per kernel (static SASS)")
print("kernel".ljust(58) + "".join(w.rjust(11) for w in watch) + "   total")
for k, c in hist.items():
    print(k[:57].ljust(58) + "".join(str(c.get(w, 0)).rjust(11) for w in watch) + f"   {sum(c.values())}")
