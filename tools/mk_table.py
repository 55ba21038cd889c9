"""Markdown rows of the throughput table from the bench lines tools/table.sh wrote:
python tools/mk_table.py [dir=profiles/r2_table] -> workload x GPU count: M segments/s, frame time, e2e, scaling efficiency T1/(N*TN)."""
import glob, json, os, re, sys
d = sys.argv[1] if len(sys.argv) > 1 else "profiles/r2_table"
rows = {}
for f in sorted(glob.glob(os.path.join(d, "r2_table_*_n*.json"))):
    m = re.match(r"r2_table_(.+)_n(\d+)\.json", os.path.basename(f))
    try:
        line = [l for l in open(f).read().splitlines() if l.strip().startswith("{")][-1]
        j = json.loads(line)
    except (IndexError, ValueError):
        continue
    rows.setdefault(m.group(1), {})[int(m.group(2))] = j
order = ["c1", "c2", "c3", "c4", "c4_group", "spheres100k", "tris1M", "c5"]
print("| workload | GPUs | M segments/s | frame | e2e M segments/s (host buffers) | efficiency T1/(N*TN) | frame CRC |")
print("|---|---:|---:|---:|---:|---:|---|")
for w in order + sorted(set(rows) - set(order)):
    if w not in rows:
        continue
    t1 = rows[w].get(1, {}).get("ms_per_step")
    for n in sorted(rows[w]):
        j = rows[w][n]
        eff = "%.3f" % (t1 / (n * j["ms_per_step"])) if t1 and n > 1 else ""
        ms = j["ms_per_step"]
        frame = "%.3f s" % (ms / 1e3) if ms >= 100 else "%.2f ms" % ms
        print("| %s | %d | %.1f | %s | %.1f | %s | %s |" % (w, n, j["value"], frame, j["e2e"]["value"], eff, j.get("frame_crc", "")))
