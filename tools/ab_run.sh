# A/B of library variants on one box: tools/ab_run.sh "" _p2 ...   (variants = suffixes of orb-slam2_b200/liborbb200<suffix>.so)
python -m pytest tests/test_gpu_extract.py tests/test_gpu_vs_reference.py -x -q > gpurun_out/ab_gputest.log 2>&1; tail -2 gpurun_out/ab_gputest.log
for v in "$@"; do
  ORBB200_LIB_PATH=$PWD/orb-slam2_b200/liborbb200$v.so python bench.py --steps 10 --warmup 3 --no-cpu --no-extras > gpurun_out/ab_bench$v.log 2>&1
  python - <<PY
import json
l=[x for x in open("gpurun_out/ab_bench$v.log") if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print("variant '$v'", d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["stage_ms_per_step"])
else:
    print(open("gpurun_out/ab_bench$v.log").read()[-1500:])
PY
done
