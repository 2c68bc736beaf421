#!/bin/bash
# Round evidence on one B200: bench lines of both arms, the ncu launch list of the bench command, full captures of the extraction
# kernels and of the small matcher kernels.  Run under gpurun from the repo root; everything lands in gpurun_out/.
R=${1:-r2}
O=gpurun_out
python bench.py --steps 10 --warmup 3 > $O/${R}_bench.log 2>&1; grep '^{' $O/${R}_bench.log | tail -1 > $O/${R}_bench_line.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/${R}_bench_ref.log 2>&1; grep '^{' $O/${R}_bench_ref.log | tail -1 > $O/${R}_bench_reference_arm_line.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $O/${R}_launches_bench_steps2.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > $O/${R}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_fast_strips|k_orient_describe|k_octree_select|k_blur7|k_stereo_match' -c 5 \
    -f -o $O/${R}_extract python bench.py --steps 1 --warmup 1 --no-cpu --no-extras > $O/${R}_ncu_extract.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_grid_build|k_window_fill|k_candidate_distances|k_triangulation_search|k_window_count|k_bf_top2" -c 14 \
    -f -o $O/${R}_match python -m pytest tests/test_gpu_match.py -x -q -k "candidate or window or triangulation or bruteforce_matches" > $O/${R}_ncu_match.log 2>&1
python tools/match_small_probe.py > $O/${R}_match_small.json 2> $O/${R}_match_small.err
ls -la $O | tail -20
tail -2 $O/${R}_bench_line.json | cut -c1-600
