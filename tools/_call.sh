mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_own_tf32_1gpu_v3.json 2> gpurun_out/bench_own_tf32_1gpu_v3.err; tail -c 300 gpurun_out/bench_own_tf32_1gpu_v3.json; echo
timeout 300 python bench.py --config mnist_fedavg --steps 5 --warmup 3 > gpurun_out/bench_mnist_fedavg_1gpu_tf32.json 2> gpurun_out/bench_mnist_tf32.err; tail -c 200 gpurun_out/bench_mnist_fedavg_1gpu_tf32.json; echo
S="eta=1,scale_weights_poison=10,lr=0.02,poison_lr=0.02,internal_epochs=4"
timeout 600 python tools/parity.py run --arm own --dtype tf32 --aggs mean --seeds 0,1,2,3,4,5,6,7,8,9,10,11 --rounds 9 --fixed-seed 1 --set $S --out gpurun_out/parity_stable_own_tf32_12seeds_fixedpart_v2.json 2> gpurun_out/parity_stable_own12f_v2.log
tail -n 3 gpurun_out/parity_stable_own12f_v2.log
timeout 200 python tools/profile_step_k.py 10 > gpurun_out/plain_k10b.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --launch-skip 262 -c 126 -o /tmp/step_k10_full -f python tools/profile_step_k.py 10 > gpurun_out/ncu_full_k10.log 2>&1
ncu -i /tmp/step_k10_full.ncu-rep --page raw --csv > gpurun_out/ncu_step_k10_raw.csv 2> gpurun_out/ncu_export.err
python tools/ncu_summary.py /tmp/step_k10_full.ncu-rep gpurun_out/ncu_step_k10_summary.txt > /dev/null 2>&1
ls -la /tmp/step_k10_full.ncu-rep gpurun_out/ncu_step_k10_raw.csv gpurun_out/ncu_step_k10_summary.txt
du -sh gpurun_out
