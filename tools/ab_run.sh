python -m pytest tests/test_gpu_extract.py tests/test_gpu_vs_reference.py tests/test_gpu_match.py -x -q > gpurun_out/r2_gputest_c.log 2>&1; tail -4 gpurun_out/r2_gputest_c.log
for v in ""; do
  ORBB200_LIB_PATH=$PWD/orb-slam2_b200/liborbb200$v.so python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2_bench_c$v.log 2>&1
  python - <<PY
import json
l=[x for x in open("gpurun_out/r2_bench_c$v.log") if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print("$v", d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["stage_ms_per_step"])
else:
    print(open("gpurun_out/r2_bench_c$v.log").read()[-1500:])
PY
done
