"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export per CUDA source line.
usage: ncu_lines.py file.csv [top]"""
import csv, sys, collections, os
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 45
cur = None; hdr = None
agg = collections.OrderedDict()
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = os.path.basename(r[1]); continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if r[0] == "" or hdr is None: continue
    d = dict(zip(hdr[4:], r[4:]))
    try:
        ie = int(d["Instructions Executed"]); te = int(d["Thread Instructions Executed"]); sm = int(d["# Samples"])
    except (KeyError, ValueError):
        continue
    k = (cur, int(r[0]))
    a = agg.setdefault(k, [0, 0, 0, r[1].strip()[:90]])
    a[0] += ie; a[1] += te; a[2] += sm
tot = [sum(a[i] for a in agg.values()) for i in range(3)]
print("total warp-inst %.3e thread-inst %.3e samples %d  avg lanes %.2f" % (tot[0], tot[1], tot[2], tot[1] / max(tot[0], 1)))
byfile = collections.Counter(); byfile_s = collections.Counter()
for (f, l), a in agg.items(): byfile[f] += a[0]; byfile_s[f] += a[2]
for f, v in byfile.most_common(): print("  %-22s inst %5.1f%%  samples %5.1f%%" % (f, 100 * v / tot[0], 100 * byfile_s[f] / max(tot[2], 1)))
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
    print("%-18s %4d inst %5.2f%% smp %5.2f%% lanes %5.1f | %s" % (f, l, 100 * a[0] / tot[0], 100 * a[2] / max(tot[2], 1), a[1] / max(a[0], 1), a[3]))
