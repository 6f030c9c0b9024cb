"""Aggregate an ncu source page per CUDA source line: python tools/ncu_source.py rep [kernel-substr] [launch-index]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]; kern = sys.argv[2] if len(sys.argv) > 2 else "k_extend"; which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
# the output is a sequence of blocks per (launch, file); split on "File Path"
blocks, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if not row: continue
    if row[0] == "File Path":
        cur = {"file": row[1], "rows": []}; blocks.append(cur)
    elif row[0] == "Function Name":
        cur["func"] = row[1]
    elif row[0] == "Line No":
        cur["hdr"] = row
    elif cur is not None and "hdr" in cur:
        cur["rows"].append(row)
# launches appear in order; group blocks by function occurrence
sel = [b for b in blocks if kern in b.get("func", "")]
files = []
seen = collections.OrderedDict()
launch = -1; last_files = set()
groups = []
for b in sel:
    if b["file"] in last_files or launch < 0:
        launch += 1; last_files = set(); groups.append([])
    last_files.add(b["file"]); groups[-1].append(b)
g = groups[which]
tot_w = tot_t = 0
lines = []
for b in g:
    h = b["hdr"]; iw = h.index("Instructions Executed"); it = h.index("Thread Instructions Executed"); ist = h.index("# Samples")
    curline = None
    for r in b["rows"]:
        if r[0] != "":
            curline = [b["file"].split("/")[-1], r[0], r[1].strip()[:90], 0, 0, 0, 0]
            lines.append(curline)
        else:
            f = lambda x: int(x) if x.strip().isdigit() else 0
            w = f(r[iw]); t = f(r[it]); s = f(r[ist])
            curline[3] += w; curline[4] += t; curline[5] += s; curline[6] += 1
            tot_w += w; tot_t += t
print("launch", which, "total warp-inst %.3e thread-inst %.3e avg thr %.1f" % (tot_w, tot_t, tot_t / max(1, tot_w)))
lines.sort(key=lambda l: -l[3])
for l in lines[:45]:
    print("%-13s %4s w=%5.1f%% thr=%4.1f samp=%5d sass=%3d | %s" % (l[0], l[1], 100 * l[3] / tot_w, l[4] / max(1, l[3]), l[5], l[6], l[2]))
