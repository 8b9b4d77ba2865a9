mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu13.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu13.log
tail -n 5 gpurun_out/pytest_gpu13.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_own_tf32_1gpu_v6.json 2> gpurun_out/bench_own_tf32_1gpu_v6.err; tail -c 600 gpurun_out/bench_own_tf32_1gpu_v6.json; echo
timeout 200 python tools/profile_step_k.py 10 > gpurun_out/plain_k10c.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel\|wgrad_tc_kernel --launch-skip 100 -c 6 -o gpurun_out/conv_wgrad_src -f python tools/profile_step_k.py 10 > gpurun_out/ncu_src.log 2>&1
ls -la gpurun_out/*.ncu-rep; du -sh gpurun_out
