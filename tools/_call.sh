mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 120 $TR --master-port 29551 bench.py --gpus 8 --steps 5 --warmup 3 --config tiny_foolsgold > gpurun_out/bench_tiny_foolsgold_8gpu.json 2> gpurun_out/bench_tiny_foolsgold_8gpu.err; tail -c 500 gpurun_out/bench_tiny_foolsgold_8gpu.json; echo; grep -v "Warning\|return func\|^\*\|OMP" gpurun_out/bench_tiny_foolsgold_8gpu.err | tail -n 5
