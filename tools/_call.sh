mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?"; tail -n 4 gpurun_out/smoke_final.log
timeout 600 python -m pytest tests/test_gpu_resnet.py -m gpu -q > gpurun_out/pytest_gpu15.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu15.log
