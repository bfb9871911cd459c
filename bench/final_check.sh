set -x
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > gpurun_out/final_gpu_tests.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -2 > gpurun_out/final_smoke.txt
timeout 200 python bench.py --impl reference --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/final_bench_ref.json
timeout 200 python bench.py --steps 20 --warmup 5 2>/dev/null | tail -1 > gpurun_out/final_bench_ours.json
cat gpurun_out/final_gpu_tests.txt gpurun_out/final_smoke.txt; cut -c1-400 gpurun_out/final_bench_ref.json; cut -c1-600 gpurun_out/final_bench_ours.json
