mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu6.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu6.log
tail -n 4 gpurun_out/pytest_gpu6.log
S="eta=1,scale_weights_poison=10,lr=0.02,poison_lr=0.02,internal_epochs=4"
timeout 600 python tools/parity.py run --arm own --dtype tf32 --aggs mean --seeds 0,1,2,3,4,5,6,7,8,9,10,11 --fixed-seed 1 --set $S --out gpurun_out/parity_stable_own_tf32_12seeds_fixedpart.json 2> gpurun_out/parity_stable_own12f.log
for c in cifar_rfa cifar_foolsgold mnist_fedavg loan_fedavg tiny_foolsgold cifar_rfa_1000; do
  timeout 900 python bench.py --config $c --steps 5 --warmup 3 > gpurun_out/bench_${c}_1gpu.json 2> gpurun_out/bench_${c}_1gpu.err; tail -c 400 gpurun_out/bench_${c}_1gpu.json; echo
done
timeout 200 python tools/profile_step_k.py 10 > gpurun_out/plain_k10.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 520 --csv --log-file gpurun_out/launches_step_k10.csv python tools/profile_step_k.py 10 > gpurun_out/ncu_k10.log 2>&1
python tools/launch_summary.py gpurun_out/launches_step_k10.csv gpurun_out/launches_step_k10.txt; head -14 gpurun_out/launches_step_k10.txt
