mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 150 $TR --master-port 29541 bench.py --gpus 8 --steps 5 --warmup 3 --config cifar_rfa_1000 > gpurun_out/bench_cifar_rfa_1000_8gpu.json 2> gpurun_out/bench_cifar_rfa_1000_8gpu.err; tail -c 500 gpurun_out/bench_cifar_rfa_1000_8gpu.json; echo; grep -v "Warning\|return func\|^\*\|OMP" gpurun_out/bench_cifar_rfa_1000_8gpu.err | tail -n 5
