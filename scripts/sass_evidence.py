"""Per-kernel SASS mnemonic evidence (tcgen05 / TMEM / TMA / vector RED / peer sync) -> profiles/sass_evidence.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "tensorflowdistributedlearning_b200/lib/libtdl_b200.so"
out = sys.argv[2] if len(sys.argv) > 2 else "profiles/sass_evidence.txt"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ["UTCHMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "UTMALDG", "UTMASTG", "UTMAPF", "SYNCS", "REDG.E.ADD.F32x4", "UBLKRED", "2CTA",
        "REDG", "LDG.E.128", "STG.E.128", "HMMA", "SHFL", "BAR.SYNC", "MEMBAR", "LD.E", "ST.E", "ATOM"]
cur, counts = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.x]+)", line)
    if m:
        op = m.group(1)
        counts[cur]["_total"] += 1
        for k in KEYS:
            if op.startswith(k) or (k == "2CTA" and ".2CTA" in op):
                counts[cur][k] += 1
lines = ["# SASS mnemonic counts per kernel of libtdl_b200.so (cuobjdump -sass), sm_100a",
         "# UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA load/store (cp.async.bulk.tensor),",
         "# UTCBAR = tcgen05.commit, UTCATOMSWS = tcgen05.alloc/dealloc, SYNCS = mbarrier ops, REDG...F32x4 = red.global.add.v4.f32",
         "# UBLKRED = cp.reduce.async.bulk (bulk reduction shared -> global), 2CTA = instructions with the cta_group::2 modifier",
         ""]
for k, c in counts.items():
    keys = [f"{n}={c[n]}" for n in KEYS if c[n]]
    lines.append(f"{k}\n    instructions={c['_total']}  " + "  ".join(keys))
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:60]))
