#!/usr/bin/env python3
"""Join an ncu `--page source --csv` SASS dump with nvdisasm -g line info:
per source line, instructions executed and stall samples.
usage: ncu_by_line.py <report.ncu-rep> <cubin> <kernel-substring> [top]"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

rep, cubin, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# collect (ordinal of instruction within the kernel) -> file:line
inside, cur, lines = False, None, []
for l in sass:
    if l.startswith("\t.section\t.text."):
        inside = kern in l
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4}\*/", l):
        lines.append((cur, l.strip()))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# the page holds one block per captured launch ("Kernel Name" row, header row, one row per instruction): take the
# first block of the wanted kernel
want = re.sub(r"I[A-Z].*$", "", kern)  # "k_render5ILi128" (mangled) -> "k_render5"
k0 = next(i for i, r in enumerate(rows) if r and r[0] == "Kernel Name" and want in r[1])
h = next(i for i in range(k0, len(rows)) if "Source" in rows[i] and "Address" in rows[i])
H = rows[h]
ii, si, sm = H.index("Instructions Executed"), H.index("Source"), H.index("# Samples")
end = next((i for i in range(h + 1, len(rows)) if rows[i] and rows[i][0] in ("Kernel Name", "Address")), len(rows))
body = rows[h + 1:end]
if len(body) != len(lines):
    print("warning: %d profiled instructions vs %d disassembled" % (len(body), len(lines)))
agg = defaultdict(lambda: [0, 0])
tot = tots = 0
for (loc, _), r in zip(lines, body):
    n, s = int(r[ii] or 0), int(r[sm] or 0)
    agg[loc][0] += n
    agg[loc][1] += s
    tot += n
    tots += s
print("total warp-instructions %d, samples %d" % (tot, tots))
for loc, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%6.2f%% inst  %6.2f%% samples  %s:%s" % (100.0 * n / tot, 100.0 * s / max(1, tots), loc[0] if loc else "?", loc[1] if loc else "?"))
