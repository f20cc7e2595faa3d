"""Per-opcode instruction / stall summary from an ncu report's source page.
   python tools/ncu_src.py rep.ncu-rep <launch index>"""
import csv, collections, subprocess, sys, io
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--launch-skip", sys.argv[2], "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1][:100])
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[hi]
si = h.index("# Samples"); src = h.index("Source"); ie = h.index("Instructions Executed")
stalls = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
seen = set(); data = []
for r in rows[hi + 1:]:
    if len(r) > si and r[si].isdigit() and r[0] not in seen:
        seen.add(r[0]); data.append(r)
agg = collections.Counter(); cnt = collections.Counter(); st = collections.defaultdict(collections.Counter); tot = collections.Counter()
for r in data:
    toks = r[src].split()
    op = toks[1] if toks[0].startswith('@') else toks[0]
    op = op.split('.')[0]
    agg[op] += int(r[si]); cnt[op] += int(r[ie])
    for i in stalls:
        if r[i].isdigit():
            st[op][h[i]] += int(r[i]); tot[h[i]] += int(r[i])
print("samples", sum(agg.values()), "warp-instr", sum(cnt.values()))
print(tot.most_common(9))
for op, n in cnt.most_common(int(sys.argv[3]) if len(sys.argv) > 3 else 18):
    print(f"{op:8s} {n:10d} samples {agg[op]:6d}", st[op].most_common(3))
