"""Count the Blackwell-specific SASS mnemonics per kernel of the built library (no GPU needed):

    python tools/sass_extract.py > profiles/r02_sass_mnemonics.csv

UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk,
UTMALDG = cp.async.bulk.tensor (tensor-map TMA), LDGSTS = cp.async; `.2CTA` marks cta_group::2;
ACQBULK / PREEXIT = griddepcontrol.wait / .launch_dependents, UCGABAR_ARV = barrier.cluster.arrive."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "umetrack_b200", "libumetrack_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTCHMMA(?:\.2CTA)?|LDTM|STTM|UTCBAR(?:\.2CTA)?(?:\.MULTICAST)?|UBLKCP|UTMALDG(?:\.\dD)?(?:\.2CTA)?|LDGSTS|HMMA|DFMA|DMUL|DADD|ACQBULK|PREEXIT|UCGABAR_ARV)\b")
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur).replace("void ", "").replace("umt::", "")
        counts[cur] = collections.Counter()
        continue
    if cur:
        for k in pat.findall(line):
            counts[cur][k] += 1
keys = sorted({k for c in counts.values() for k in c})
w = sys.stdout
w.write("kernel," + ",".join(keys) + "\n")
for name, c in counts.items():
    if sum(c.values()):
        w.write('"' + name + '",' + ",".join(str(c.get(k, 0)) for k in keys) + "\n")
