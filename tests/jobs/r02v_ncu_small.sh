# round-2: ncu --set full of the three instruction-bound "HBM" kernels (sample_coarse, sample_fine, point_taps) at 2^20 rays
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --kernels"
$CMD > gpurun_out/r02v_plain.json 2> gpurun_out/r02v_plain.err; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sample_coarse4_kernel|sample_fine_kernel|point_features_kernel" -s 6 -c 60 -f -o gpurun_out/r02v_small $CMD > gpurun_out/r02v_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r02v_ncu.log
ncu -i gpurun_out/r02v_small.ncu-rep --page raw --csv > gpurun_out/r02v_small_raw.csv 2>/dev/null; echo "export rc=$?"
rm -f gpurun_out/r02v_small.ncu-rep
python - <<'PY'
import csv, json
rows = list(csv.DictReader(open("gpurun_out/r02v_small_raw.csv")))
want = ["Kernel Name", "gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum", "smsp__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.avg.per_cycle_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio" ]
seen = {}
for r in rows[1:]:
    name = r["Kernel Name"].split("(")[0]
    seen[name] = {k: r.get(k) for k in want if k in r}
json.dump(seen, open("gpurun_out/r02v_small_summary.json", "w"), indent=1)
for k, v in seen.items():
    print("NCU", k[:50], {a.split(".")[0][:28]: b for a, b in v.items() if a != "Kernel Name"})
PY
