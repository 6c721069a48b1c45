"""Developer tool: join an ncu SASS source page (--page source --csv) with nvdisasm -g line
info of the same cubin and print instructions executed / stall samples per source line.
usage: ncu_lines.py <src.csv> <disasm.sass> <mangled-kernel-substring> [top]"""
import csv, re, sys, collections
src_csv, sass, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
# nvdisasm: offset -> (file, line) innermost + inline chain
lines = open(sass).read().split("\n")
inside = False
cur = None
off2line = {}
for ln in lines:
    if ln.startswith("//---") and ".text." in ln:
        inside = kern in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)), m.group(3).strip())
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*)", ln)
    if m:
        off2line[int(m.group(1), 16)] = (cur, m.group(2).strip())
rows = list(csv.reader(open(src_csv)))
hdr = next(r for r in rows if r and r[0] == "Address")
data = [dict(zip(hdr, r)) for r in rows if len(r) == len(hdr) and r[0] != "Address"]
a0 = int(data[0]["Address"], 16) if data[0]["Address"].startswith("0x") else int(data[0]["Address"])
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for d in data:
    a = int(d["Address"], 16) if d["Address"].startswith("0x") else int(d["Address"])
    key, ins = off2line.get(a - a0, (None, "?"))
    ie = int(d["Instructions Executed"] or 0)
    te = int(d["Thread Instructions Executed"] or 0)
    sm = int(d["# Samples"] or 0)
    k = key[:2] if key else ("?", 0)
    for i, v in enumerate((ie, te, sm)):
        agg[k][i] += v
        tot[i] += v
print("total warp-instr %d  thread-instr %d  samples %d" % tuple(tot))
srcs = {}
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    if f not in srcs:
        try:
            import glob
            p = glob.glob("/root/repo/**/" + f, recursive=True)[0]
            srcs[f] = open(p).read().split("\n")
        except Exception:
            srcs[f] = []
    text = srcs[f][l - 1].strip()[:90] if 0 < l <= len(srcs[f]) else ""
    print("%5.1f%% instr %5.1f%% smp  lanes %4.1f  %s:%d  %s" % (100 * v[0] / tot[0], 100 * v[2] / max(tot[2], 1), v[1] / max(v[0], 1), f, l, text))
if len(sys.argv) > 5:
    # per-address dump for one source line: clusters of inlined copies
    f, l = sys.argv[5].split(":")
    for d in data:
        a = int(d["Address"], 16) if d["Address"].startswith("0x") else int(d["Address"])
        key, ins = off2line.get(a - a0, (None, "?"))
        if key and key[0] == f and key[1] == int(l):
            print("%06x %10s %10s  %s" % (a - a0, d["Instructions Executed"], d["Thread Instructions Executed"], ins[:60]))
