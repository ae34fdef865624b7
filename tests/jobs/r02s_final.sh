# round-2 final single-GPU evidence: whole GPU suite, the no-alloc test alone afterwards, bench line, reference arm
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02s_gpu_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/r02s_gpu_tests.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02s_noalloc_alone.log 2>&1; echo "noalloc rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02s_smoke.log 2>&1; echo "smoke rc=$?"
tail -2 gpurun_out/r02s_smoke.log
timeout 900 python bench.py > gpurun_out/r02s_bench_n1.json 2> gpurun_out/r02s_bench_n1.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r02s_bench_n1.json").read().strip().splitlines()[-1])
print("BENCH", round(d["value"]), round(d["e2e"]["value"]), round(d["ms_per_step"], 2), round(d["roofline"]["frac"], 3), d["clocks"], d["gpu_launches"])
print("SHARDED", {k: (round(v.get("rays/s", 0)), round(v.get("ms_per_call", 0), 1)) for k, v in d.get("sharded", {}).items() if isinstance(v, dict)})
PY
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02s_bench_reference.json 2> gpurun_out/r02s_bench_reference.err; echo "ref rc=$?"
tail -c 400 gpurun_out/r02s_bench_reference.json
