# round-2 final launch list: ncu gpu__time_duration of the bench command (after the same command exited 0 without ncu)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-extras --no-cpu-baseline"
$CMD > gpurun_out/r02u_plain.json 2> gpurun_out/r02u_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02u_launches_bench.csv $CMD > gpurun_out/r02u_ncu.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import csv, collections, json
rows = []
with open("gpurun_out/r02u_launches_bench.csv") as f:
    lines = [l for l in f if not l.startswith("==")]
r = csv.DictReader(lines)
tot = collections.defaultdict(lambda: [0, 0.0])
for row in r:
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = row["Kernel Name"].split("(")[0]
    v = float(row["Metric Value"].replace(",", ""))
    unit = row["Metric Unit"]
    us = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
    tot[name][0] += 1
    tot[name][1] += us
s = sum(v[1] for v in tot.values())
out = {"command": "ncu --metrics gpu__time_duration.sum --clock-control none -c 400 python bench.py --steps 2 --warmup 1 --no-extras --no-cpu-baseline",
       "total_us": s, "kernels": {k: {"launches": v[0], "us": round(v[1], 1), "share": round(v[1] / s, 4)} for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1])}}
json.dump(out, open("gpurun_out/r02u_launches_bench_summary.json", "w"), indent=1)
for k, v in list(out["kernels"].items())[:8]:
    print("LAUNCH", k[:60], v)
PY
