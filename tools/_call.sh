mkdir -p gpurun_out
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_reference_1gpu.json 2> gpurun_out/bench_reference_1gpu.err; tail -c 500 gpurun_out/bench_reference_1gpu.json; echo
