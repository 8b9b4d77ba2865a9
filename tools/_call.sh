mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu7.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu7.log
tail -n 12 gpurun_out/pytest_gpu7.log
timeout 300 python tools/gpu_step_cost.py > gpurun_out/step_cost_stage_pack.txt 2>&1; cat gpurun_out/step_cost_stage_pack.txt
timeout 300 python tools/gpu_step_cost.py mnist > gpurun_out/step_cost_mnist_tf32.txt 2>&1; cat gpurun_out/step_cost_mnist_tf32.txt
timeout 300 python bench.py --config mnist_fedavg --steps 5 --warmup 3 > gpurun_out/bench_mnist_fedavg_1gpu_tf32.json 2> gpurun_out/bench_mnist_tf32.err; tail -c 300 gpurun_out/bench_mnist_fedavg_1gpu_tf32.json
S="eta=1,scale_weights_poison=10,lr=0.02,poison_lr=0.02,internal_epochs=4"
timeout 900 python tools/parity.py run --arm own --dtype tf32 --aggs mean --seeds 0,1,2,3,4,5 --rounds 9 --fixed-seed 1 --set $S,backend=torch --out gpurun_out/parity_stable_own_torchbackend.json 2> gpurun_out/parity_stable_torchbackend.log
tail -n 7 gpurun_out/parity_stable_torchbackend.log
