#!/usr/bin/env python3
"""Per-source-line instruction totals of the render kernel: joins the per-SASS 'Instructions Executed'
column of an .ncu-rep (source page) with nvdisasm's line info of the SAME librt_b200.so build.

    python tools/sass_lines.py gpurun_out/prof.ncu-rep [lib.so] [kernel-substring] [out.txt]
"""
import csv, io, os, re, subprocess, sys, tempfile

rep = sys.argv[1]
lib = sys.argv[2] if len(sys.argv) > 2 else "opengl-ray-tracer_b200/csrc/librt_b200.so"
kern = sys.argv[3] if len(sys.argv) > 3 else "render_kernelILb0ELb0ELi8E"
outp = sys.argv[4] if len(sys.argv) > 4 else None

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.split("\n")
lines, cur, inside = [], ("?", 0), False          # (file, line) per SASS instruction of the kernel, in order
for ln in dis:
    if ln.startswith("\t.section") or ln.startswith(".section"):
        inside = (".text." in ln) and (kern in ln)
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        lines.append(cur)

raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]; data = rows[2:]
ie = hdr.index("Instructions Executed"); ism = hdr.index("Warp Stall Sampling (All Samples)")
if len(data) != len(lines):
    print("WARNING: %d SASS rows in the report vs %d in the library (different build?)" % (len(data), len(lines)))
n = min(len(data), len(lines))
tot = sum(int(r[ie]) for r in data); tots = sum(int(r[ism]) for r in data)
agg = {}
for i in range(n):
    a = agg.setdefault(lines[i], [0, 0, 0])
    a[0] += int(data[i][ie]); a[1] += int(data[i][ism]); a[2] += 1
src = {}
def srcline(f, l):
    if f not in src:
        for d in ("opengl-ray-tracer_b200/csrc", "include"):
            p = os.path.join(d, f)
            if os.path.exists(p):
                src[f] = open(p).read().split("\n"); break
        else:
            src[f] = []
    return src[f][l - 1].strip()[:90] if 0 < l <= len(src[f]) else ""
out = ["# total warp instructions %d, stall samples %d, SASS instructions %d" % (tot, tots, n),
       "# %inst  %samples  nSASS  file:line  source"]
for (f, l), (e, s, c) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:70]:
    out.append("%5.1f%% %5.1f%% %4d  %s:%d  %s" % (100.0 * e / tot, 100.0 * s / max(tots, 1), c, f, l, srcline(f, l)))
txt = "\n".join(out)
print(txt)
if outp:
    open(outp, "w").write(txt + "\n")
