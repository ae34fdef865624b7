# round-2 GPU job: smoke() as the driver runs it, then the CPU-side contract of the bench's reference arm
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02n_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/r02n_smoke.log
tail -6 gpurun_out/r02n_smoke.log | cut -c1-300
timeout 600 python bench.py --train --train-rays 128 4096 --steps 5 > gpurun_out/r02n_train.json 2> gpurun_out/r02n_train.err; echo "train rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02n_train.json").read().strip().splitlines()[-1])
for r in d["sweep"]: print("TRAIN", r["rays_per_step"], "rays", round(r["ms_per_step"], 2), "ms", round(r["rays/s"]), "rays/s", r["kernel_ms"])
PY
