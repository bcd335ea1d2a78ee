"""Aggregate ncu warp-stall samples per CUDA source line.
usage: python scripts/ncu_lines.py <report.ncu-rep> <lib.so> <kernel-regex> [top]
Correlates the SASS page of the report (one row per instruction, in address order) with
`nvdisasm --print-line-info` of the same kernel extracted from the .so."""
import csv, os, re, subprocess, sys, tempfile, collections

rep, so, kre = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# first launch only
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur); continue
    if cur is None: continue
    if r and r[0] == "Address": cur["hdr"] = r; continue
    if len(r) > 5: cur["rows"].append(r)
b = blocks[0]
h = b["hdr"]; iS = h.index("# Samples"); iI = h.index("Instructions Executed"); iT = h.index("Thread Instructions Executed")
mangled = None
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
# split per function
fn = None; lines_by_fn = collections.defaultdict(list); curline = ("?", 0)
for ln in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m: fn = m.group(1); curline = ("?", 0); continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: curline = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if fn and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        lines_by_fn[fn].append(curline)
short = re.sub(r"\(.*", "", b["name"])
cands = [f for f in lines_by_fn if short in f]
assert cands, (short, list(lines_by_fn)[:5])
cands.sort(key=lambda f: abs(len(lines_by_fn[f]) - len(b["rows"])))
li = lines_by_fn[cands[0]]
n = min(len(li), len(b["rows"]))
print(f"# {b['name'][:60]}  sass rows {len(b['rows'])}  disasm instrs {len(li)}")
agg = collections.Counter(); inst = collections.Counter(); tin = collections.Counter()
for k in range(n):
    r = b["rows"][k]
    agg[li[k]] += int(r[iS]); inst[li[k]] += int(r[iI]); tin[li[k]] += int(r[iT])
tot = sum(agg.values()); ti = sum(inst.values())
print(f"# total samples {tot}, warp instr {ti}, thread instr {sum(tin.values())}")
byfile = collections.Counter()
for (f, l), s in agg.items(): byfile[f] += s
print("# by file:", dict(byfile))
for (f, l), s in agg.most_common(top):
    print(f"{100*s/tot:5.1f}%  inst {100*inst[(f,l)]/ti:5.1f}%  lanes {tin[(f,l)]/max(1,inst[(f,l)]):4.1f}  {f}:{l}")
