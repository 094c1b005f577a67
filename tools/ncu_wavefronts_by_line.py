#!/usr/bin/env python3
"""Per source line: shared-memory wavefronts (actual / ideal), global sectors and
instructions, from an ncu report joined with nvdisasm -g line info.
usage: ncu_wavefronts_by_line.py <report.ncu-rep> <cubin> <kernel-substring> [top]"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

rep, cubin, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 16
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
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
        lines.append(cur)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = next(i for i, r in enumerate(rows) if "Source" in r and "Address" in r)
H, body = rows[h], rows[h + 1:]
cols = {k: H.index(k) for k in ("L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal", "Instructions Executed", "L2 Theoretical Sectors Global")}
agg = defaultdict(lambda: [0, 0, 0, 0])
for loc, r in zip(lines, body):
    for j, k in enumerate(cols):
        try:
            agg[loc][j] += int(r[cols[k]] or 0)
        except ValueError:
            pass
tot = [sum(v[j] for v in agg.values()) for j in range(4)]
print("total: smem wavefronts %d (ideal %d), instructions %d, global sectors %d" % tuple(tot))
for loc, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%6.2f%% wf=%d ideal=%d inst=%d gsect=%d %s:%s" % (100.0 * v[0] / max(1, tot[0]), v[0], v[1], v[2], v[3], loc[0] if loc else "?", loc[1] if loc else "?"))
