mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu11.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu11.log
tail -n 4 gpurun_out/pytest_gpu11.log
timeout 300 python tools/gpu_step_cost.py > gpurun_out/step_cost_wgrad2p.txt 2>&1; cat gpurun_out/step_cost_wgrad2p.txt
