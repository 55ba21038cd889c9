#!/bin/bash
# does ray coherence help the shared-memory-tree walk (config 4 geometry)?  per-kernel durations with and without sorted waves
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
for so in 0 1; do
  SRT_WF_SORT=$so timeout 600 python tools/prof_one.py c3_masterchief 1280 720 40 4 2 > $O/al_run_$so.log 2>&1; cat $O/al_run_$so.log
  SRT_WF_SORT=$so timeout 900 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum --clock-control none -c 600 --csv --log-file $O/al_launches_$so.csv \
    python tools/prof_one.py c3_masterchief 1280 720 40 4 2 > /dev/null 2>&1
done
python - <<'P'
import csv, collections
for so in (0, 1):
    rows = list(csv.reader(open('gpurun_out/al_launches_%d.csv' % so)))
    hdr = None; agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for r in rows:
        if hdr is None:
            if 'Kernel Name' in r: hdr = r
            continue
        d = dict(zip(hdr, r))
        k = d['Kernel Name'].split('(')[0][:50]
        v = float(d['Metric Value'].replace(',', ''))
        if d['Metric Name'] == 'gpu__time_duration.sum':
            agg[k][0] += 1; agg[k][1] += v * {'ns': 1e-6, 'us': 1e-3, 'ms': 1}.get(d['Metric Unit'], 1e-6)
        elif d['Metric Name'].startswith('smsp__thread_inst'): agg[k][2] += v
        elif d['Metric Name'] == 'smsp__inst_executed.sum': agg[k][3] += v
    print("== SRT_WF_SORT=%d" % so)
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1])[:8]:
        print("  %-52s %4d launches %9.2f ms  lanes %.1f  warp-inst %.3e" % (k, v[0], v[1], v[2] / max(v[0], 1), v[3]))
P
