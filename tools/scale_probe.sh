TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 1 4 8; do
  timeout 120 $TR --nproc-per-node $n --master-port $((29600+n)) tools/h2d_probe.py 2>/dev/null | grep '^{' > gpurun_out/r2_h2d_n${n}.json
  timeout 120 $TR --nproc-per-node $n --master-port $((29700+n)) tools/h2d_probe.py --bind 2>/dev/null | grep '^{' > gpurun_out/r2_h2d_n${n}_bind.json
done
timeout 300 $TR --nproc-per-node 8 --master-port 29801 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu --no-extras 2>/dev/null | grep '^{' > gpurun_out/r2_bench_8gpu_bind.json
timeout 300 $TR --nproc-per-node 8 --master-port 29802 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu --no-extras --no-bind 2>/dev/null | grep '^{' > gpurun_out/r2_bench_8gpu_nobind.json
lscpu | grep -E "Model name|Socket|NUMA|^CPU\(s\)" > gpurun_out/r2_lscpu.txt
nvidia-smi topo -m > gpurun_out/r2_topo.txt 2>&1
ls -la gpurun_out/r2_h2d* gpurun_out/r2_bench_8gpu*
