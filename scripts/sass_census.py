"""One-screen opcode census per kernel of libskp_b200.so (cuobjdump -sass): the SASS mnemonics that prove a Blackwell-native
path (B200_PROFILING.md): UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG = TMA loads, UTCBAR = tcgen05.commit,
SYNCS = mbarrier, MUFU.EX2, REDG = red.global, USETMAXREG = setmaxnreg, HMMA would be the legacy mma.sync path (none expected).
python scripts/sass_census.py > profiles/r2_sass_census.txt"""
import collections, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(root, "stablekeypoints_b200", "libskp_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
OPS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMAREDG", "SYNCS", "MUFU.EX2", "REDG", "USETMAXREG", "HMMA", "UBLKCP", "FFMA", "LDS", "STS", "LDG", "STG", "ATOMG"]
cur, counts, total = None, collections.OrderedDict(), collections.Counter()
for ln in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        cur = cur.replace("void ", "").replace("skp::", "")
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
    if m:
        op = m.group(1)
        counts[cur]["_all"] += 1
        for o in OPS:
            if op.startswith(o):
                counts[cur][o] += 1
                total[o] += 1
print(f"# {os.path.relpath(lib, root)}: SASS opcode census per kernel (cuobjdump -sass, sm_100a)")
print("# " + " ".join(f"{o:>9s}" for o in ["instrs"] + OPS[:11]) + "  kernel")
for k, c in counts.items():
    if c["_all"] == 0:
        continue
    print("  " + " ".join(f"{c[o]:9d}" for o in ["_all"] + OPS[:11]) + "  " + k)
print("# totals: " + ", ".join(f"{o} {total[o]}" for o in OPS[:11]))
