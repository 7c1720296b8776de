#!/usr/bin/env python
"""Per-source-line instruction counts of one kernel launch in an ncu report.
Joins `ncu --page source --print-source sass` (per-SASS-instruction executed counts) with `nvdisasm -gi` of the cubin
inside libolrt.so (SASS offset -> file:line, with the inlined-at chain).
usage: ncu_lines.py report.ncu-rep launch_index mangled_kernel_substring nrays [top]"""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, launch, kname, nrays = sys.argv[1], int(sys.argv[2]), sys.argv[3], int(sys.argv[4])
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
so = os.path.join(ROOT, "optiland_pr_b200", "libolrt.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
sass = ""
for f in os.listdir(tmp):
    if f.endswith(".cubin"):
        out = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        if kname in out:
            sass = out
            break
sec = None
lines = {}     # offset -> (innermost "file:line", outermost "file:line")
cur = None
fresh = True
for ln in sass.split("\n"):
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
    if m:
        sec = m.group(1)
        continue
    if sec is None or kname not in sec:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:   # consecutive comment lines: innermost frame first, then the inlined-at chain outward
        here = os.path.basename(m.group(1)) + ":" + m.group(2)
        chain = re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3))
        outer = (os.path.basename(chain[-1][0]) + ":" + chain[-1][1]) if chain else here
        if fresh:
            cur = (here, outer)
            fresh = False
        else:
            cur = (cur[0], outer)
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m and cur is not None:
        lines[int(m.group(1), 16)] = (cur, m.group(2).strip())
        fresh = True
src = subprocess.run(["ncu", "-i", rep, "--launch-skip", str(launch), "--launch-count", "1", "--page", "source", "--csv",
                      "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = None
body = []
for r in rows:
    if r and r[0] == "Address":
        hdr = r
        continue
    if hdr and len(r) > 6:
        body.append(r)
iE, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
base = int(body[0][0], 16)
nw = nrays / 32
by_inner, by_outer, samp_inner = collections.Counter(), collections.Counter(), collections.Counter()
ops_inner = collections.defaultdict(collections.Counter)
tot = 0
seen = set()
for r in body:
    off = int(r[0], 16) - base
    if off in seen:
        continue
    seen.add(off)
    n = int(r[iE])
    tot += n
    (inner, outer), text = lines.get(off, (("?", "?"), r[1]))
    by_inner[inner] += n
    by_outer[outer] += n
    samp_inner[inner] += int(r[iS])
    t = r[1].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    ops_inner[inner][op] += n
print(f"total warp-instructions per ray: {tot / nw:.0f}")
print("--- by innermost source line (instr/ray, samples, top opcodes)")
for k, n in by_inner.most_common(top):
    print(f"{k:28s} {n / nw:8.1f} {samp_inner[k]:7d}  " + " ".join(f"{o}:{c / nw:.0f}" for o, c in ops_inner[k].most_common(5)))
print("--- by line of the kernel body (inlined callees folded in)")
for k, n in by_outer.most_common(top):
    print(f"{k:28s} {n / nw:8.1f}")
