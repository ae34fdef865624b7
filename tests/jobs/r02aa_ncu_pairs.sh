# round-2 ncu job (ONE ncu use per gpurun call): --set full of the chain launches of one precise-chain frame.
#   PREC=fp16x2|fp16x3 bash tests/jobs/r02aa_ncu_pairs.sh
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
PREC=${PREC:-fp16x2}
CMD="python tests/ncu_frame.py --scene srn_car --frames 2 --precision $PREC"
$CMD > gpurun_out/r02aa_plain_$PREC.log 2>&1; echo "plain rc=$?"; tail -1 gpurun_out/r02aa_plain_$PREC.log
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:mlp_chain_kernel -s 2 -c 2 -f -o gpurun_out/r02aa_chain_$PREC $CMD > gpurun_out/r02aa_ncu_$PREC.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r02aa_chain_$PREC.ncu-rep --page raw --csv > gpurun_out/r02aa_chain_${PREC}_raw.csv 2>/dev/null; echo "export rc=$?"
rm -f gpurun_out/r02aa_chain_$PREC.ncu-rep
python - <<PY
import csv, json
rows = list(csv.DictReader(open("gpurun_out/r02aa_chain_${PREC}_raw.csv")))
keys = ["gpu__time_duration.sum", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_tensor.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__cycles_elapsed.avg.per_second",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct"]
out = []
for r in rows[1:]:
    out.append({"kernel": r["Kernel Name"][:60], **{k: r.get(k) for k in keys if r.get(k) not in (None, "")}})
    tens = [k for k in r if "tensor" in k and "pct" in k]
    out[-1]["tensor_pct_metrics"] = {k: r[k] for k in tens[:8]}
json.dump(out, open("gpurun_out/r02aa_chain_${PREC}_summary.json", "w"), indent=1)
for o in out:
    print("NCU", {k: v for k, v in o.items() if k != "tensor_pct_metrics"})
    print("NCUT", o["tensor_pct_metrics"])
PY
