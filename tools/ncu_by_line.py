#!/usr/bin/env python
"""Join an ncu SASS source page with nvdisasm line info: executed warp instructions and stall samples per source line.

usage: tools/ncu_by_line.py <report.ncu-rep> <function-name>[@<substring of the demangled name>] [<mangled-substring>] [top]
Needs ncu, cuobjdump and nvdisasm (all present in the build container; no GPU needed)."""
import collections
import csv
import glob
import io
import os
import re
import subprocess
import sys
import tempfile

rep, kname = sys.argv[1], sys.argv[2]
mangled = sys.argv[3] if len(sys.argv) > 3 else kname
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "openhyperflow2d_b200", "libhf2d_gpu.so")

fn, _, sub = kname.partition("@")
if rep.endswith(".csv.gz"):      # a source page exported on the GPU box (tools/prof_plain.sh): every captured kernel, one block each
    import gzip
    out = gzip.open(rep, "rt").read()
else:
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + fn], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# several launches may follow each other: take the first block whose demangled name holds `sub`
k_i = next(i for i, r in enumerate(rows) if r and r[0] == "Kernel Name" and sub in r[1] and (fn + "<") in r[1].replace(" ", "") + "<")
hdr_i = next(i for i, r in enumerate(rows) if i > k_i and r and r[0] == "Address")
hdr = rows[hdr_i]
body = []
for r in rows[hdr_i + 1:]:
    if not r or r[0] in ("Kernel Name", "Address"):
        break
    body.append(r)
ci = {h: i for i, h in enumerate(hdr)}

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
cub = glob.glob(os.path.join(tmp, "*.cubin"))[0]
dis = subprocess.run(["nvdisasm", "--print-line-info", cub], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and mangled in l)
lines = []
cur = ("?", 0)
for l in dis[start + 1:]:
    if l.startswith(".text.") or l.startswith("\t.section") or l.startswith(".section"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        lines.append((cur, m.group(2).strip()))
if len(lines) != len(body):
    print(f"warning: {len(lines)} disassembled instructions vs {len(body)} profiled", file=sys.stderr)
n = min(len(lines), len(body))
by_line = collections.Counter(); samp = collections.Counter(); by_op = collections.Counter(); by_file = collections.Counter()
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
by_reason = collections.Counter(); line_reason = collections.defaultdict(collections.Counter)
tot = 0
for (loc, ins), r in zip(lines[:n], body[:n]):
    ex = int(r[ci["Instructions Executed"]] or 0)
    sm = int(r[ci["# Samples"]] or 0)
    by_line[loc] += ex; samp[loc] += sm; by_file[loc[0]] += ex
    for h in stall_cols:
        v = int(r[ci[h]] or 0)
        if v:
            by_reason[h] += v; line_reason[loc][h] += v
    op = re.sub(r"^@!?U?P\d+\s+", "", ins).split()[0].split(".")[0]
    by_op[op] += ex
    tot += ex
print(f"total warp instructions executed: {tot}")
print("-- by file"); [print(f"{v:12d} {100 * v / tot:5.1f}%  {k}") for k, v in by_file.most_common()]
print("-- by opcode"); [print(f"{v:12d} {100 * v / tot:5.1f}%  {k}") for k, v in by_op.most_common(25)]
print("-- by line (instructions, samples)")
ts = sum(samp.values()) or 1
for k, v in by_line.most_common(top):
    print(f"{v:12d} {100 * v / tot:5.1f}%  samples {100 * samp[k] / ts:5.1f}%  {k[0]}:{k[1]}")
print("-- stall samples by reason")
tr = sum(by_reason.values()) or 1
for k, v in by_reason.most_common(10):
    print(f"{100 * v / tr:5.1f}%  {k}")
print("-- lines by stall samples (dominant reasons)")
for k, v in samp.most_common(top):
    rs = ", ".join(f"{a.replace('stall_', '')} {100 * b / max(v, 1):.0f}%" for a, b in line_reason[k].most_common(3))
    print(f"samples {100 * v / ts:5.1f}%  instr {100 * by_line[k] / tot:5.1f}%  {k[0]}:{k[1]}  [{rs}]")
