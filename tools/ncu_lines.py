"""Dynamic instruction counts of one profiled kernel, by opcode and by source line (development aid).
usage: python tools/ncu_lines.py report.ncu-rep <mangled-kernel-substring> [top N]
Joins ncu's SASS page (per-instruction executed counts) with nvdisasm -g line markers of
fo-rma_b200/csrc/build/kernels.cu.o by instruction offset."""
import collections, csv, glob, io, os, re, subprocess, sys, tempfile
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
obj = os.path.join(ROOT, "fo-rma_b200", "csrc", "build", os.environ.get("NCU_LINES_OBJ", "kernels.cu.o"))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=tmp, capture_output=True)
dis = subprocess.run(["nvdisasm", "-g", "-c", glob.glob(tmp + "/*.cubin")[0]], capture_output=True, text=True).stdout
line_of, cur, inside = {}, ("?", 0), False
for l in dis.splitlines():
    if l.startswith("//---") and ".text." in l: inside = kern in l
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", l)
    if m: line_of[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
ci = {k: i for i, k in enumerate(rows[hdr])}
base = None
ops, opst, lines, linest, stall = (collections.Counter() for _ in range(5))
ts = 0
tw = tt = 0
for r in rows[hdr + 1:]:
    try:
        addr = int(r[0], 16); w = int(r[ci["Instructions Executed"]]); t = int(r[ci["Thread Instructions Executed"]])
    except (ValueError, IndexError):
        continue
    if base is None: base = addr
    tw += w; tt += t
    op = next((x for x in r[ci["Source"]].split() if not x.startswith("@")), "?").split(".")[0]
    ops[op] += w; opst[op] += t
    k = line_of.get(addr - base, ("?", 0)); lines[k] += w; linest[k] += t
    try: sm = int(r[ci["Warp Stall Sampling (All Samples)"]])
    except (ValueError, KeyError): sm = 0
    stall[k] += sm; ts += sm
print(f"warp instr {tw:.4g}  thread instr {tt:.4g}  avg active {tt / max(tw, 1):.2f}")
for op, w in ops.most_common(top // 2):
    print(f"{op:10s} {w:14d} {100 * w / tw:6.2f}%  active {opst[op] / max(w, 1):5.1f}")
print("---- by source line")
src = {}
for (f, n), w in lines.most_common(top):
    if f not in src:
        p = os.path.join(ROOT, "fo-rma_b200", "csrc", f)
        src[f] = open(p).read().splitlines() if os.path.exists(p) else []
    text = src[f][n - 1].strip()[:90] if 0 < n <= len(src[f]) else ""
    print(f"{f}:{n:<4d} {100 * w / tw:6.2f}%  active {linest[(f, n)] / max(w, 1):5.1f}  stall {100 * stall[(f, n)] / max(ts, 1):5.2f}%  {text}")
print("---- by stall samples")
for (f, n), sm in stall.most_common(top // 2):
    text = src.get(f, [])[n - 1].strip()[:90] if f in src and 0 < n <= len(src[f]) else ""
    print(f"{f}:{n:<4d} stall {100 * sm / max(ts, 1):5.2f}%  instr {100 * lines[(f, n)] / tw:5.2f}%  {text}")
