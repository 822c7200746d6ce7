"""SASS evidence of libspanon.so: mnemonic counts per kernel for the tcgen05 / TMEM / bulk-copy / cluster / PDL paths.

    python tools/sass_summary.py > profiles/r02_sass_tcgen05.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "speech_anonymization_b200", "libspanon.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
KEYS = ["UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "SYNCS", "UCGABAR", "HMMA", "MUFU.TANH", "FFMA2", "FADD2", "ACQBULK", "PREEXIT"]
print("# SASS evidence of libspanon.so (cuobjdump -sass, sm_100a): mnemonic counts per kernel for the tcgen05 / TMEM / bulk-copy / cluster / PDL paths")
print("# UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / .st, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk (1-D bulk copy: the chunk-planar layout makes")
print("# every operand slab one contiguous range, so no tensor-map TMA / UTMALDG is used), SYNCS = mbarrier ops, HMMA = mma.sync, UCGABAR = barrier.cluster,")
print("# FFMA2 / FADD2 = packed fp32 pairs, ACQBULK / PREEXIT = griddepcontrol.wait / .launch_dependents (programmatic dependent launch)\n")
name, counts, n = None, collections.Counter(), 0


def flush():
    if name is None:
        return
    try:
        dem = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
    except OSError:
        dem = name
    dem = re.sub(r"^void ", "", dem)
    dem = re.sub(r"spanon::|\(anonymous namespace\)::|<unnamed>::", "", dem)
    dem = dem.replace("(int)", "").replace("(bool)", "")
    short = re.sub(r"\(.*$", "", dem)          # drop the argument list
    hits = " ".join("%s=%d" % (k, counts[k]) for k in KEYS if counts[k])
    if hits:
        print("%-46s insts %5d  %s" % (short[:46], n, hits))


for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        flush()
        name, counts, n = m.group(1), collections.Counter(), 0
        continue
    m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        n += 1
        op = m.group(1)
        for k in KEYS:
            if op == k or op.startswith(k + "."):
                counts[k] += 1
flush()
