"""SASS opcode counts per kernel of the shipped library -> profiles/r02_sass_opcodes.tsv
    python tools/sass_opcodes.py > profiles/r02_sass_opcodes.tsv"""
import collections
import os
import re
import subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dist-dqn_b200", "libdqn_b200.so")
OPS = ["UTCHMMA", "LDTM", "UBLKCP", "UTMALDG", "UTCBAR", "SYNCS", "LDGSTS", "UCGABAR_ARV", "HMMA", "FFMA"]
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
names = {}
counts = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1)
        counts[cur]["instructions"] += 1
        for o in OPS:
            if op == o or op.startswith(o + "."):
                counts[cur][o] += 1
dem = subprocess.run(["cu++filt"] + list(counts), capture_output=True, text=True).stdout.splitlines()
print("# SASS opcode counts per kernel of dist-dqn_b200/libdqn_b200.so (cuobjdump -sass, built for sm_100a); round 2")
print("# UTCHMMA = tcgen05.mma | LDTM = tcgen05.ld | UBLKCP = cp.async.bulk (TMA engine) | UTCBAR = tcgen05.commit | SYNCS = mbarrier ops")
print("# LDGSTS = cp.async | UCGABAR_ARV = barrier.cluster.arrive (thread-block cluster) | HMMA = legacy mma.sync (none: every tensor-core kernel is tcgen05)")
print("\t".join(["kernel", "instructions"] + OPS))
for (k, c), d in zip(counts.items(), dem):
    short = re.sub(r"\((?:[^()]|\([^()]*\))*\)$", "", d)          # drop the parameter list, keep template arguments
    short = re.sub(r"^void ", "", short)
    print("\t".join([short[:150], str(c["instructions"])] + [str(c[o]) for o in OPS]))
