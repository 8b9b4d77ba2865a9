mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu4.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu4.log
tail -n 8 gpurun_out/pytest_gpu4.log
timeout 300 python tools/gpu_step_cost.py > gpurun_out/step_cost_bnsums.txt 2>&1; cat gpurun_out/step_cost_bnsums.txt
DBA_FUSE_BN_SUMS=0 timeout 300 python tools/gpu_step_cost.py > gpurun_out/step_cost_nobnsums.txt 2>&1; cat gpurun_out/step_cost_nobnsums.txt
timeout 600 python tools/gpu_mnist_compare.py > gpurun_out/mnist_compare.txt 2>&1; cat gpurun_out/mnist_compare.txt | tail -8
S="eta=1,scale_weights_poison=10,lr=0.02,poison_lr=0.02,internal_epochs=4"
timeout 600 python tools/parity.py run --arm own --dtype tf32 --aggs mean --seeds 0,1,2,3,4,5,6,7,8,9,10,11 --set $S --out gpurun_out/parity_stable_own_tf32_12seeds.json 2> gpurun_out/parity_stable_own12.log
timeout 200 python tools/profile_step_k.py 1 > gpurun_out/plain_k1.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 450 --csv --log-file gpurun_out/launches_step_k1_v2.csv python tools/profile_step_k.py 1 > gpurun_out/ncu_k1.log 2>&1
python tools/launch_summary.py gpurun_out/launches_step_k1_v2.csv gpurun_out/launches_step_k1_v2.txt; head -12 gpurun_out/launches_step_k1_v2.txt
