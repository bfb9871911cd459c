# A/B of programmatic dependent launch (PDNN_PDL=1) on one GPU: numerics first, then the flagship step both ways.
PDNN_PDL=1 timeout 400 python -m pytest tests/test_gpu_engine.py tests/test_gpu_kernels.py tests/test_gpu_canary.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -3 > gpurun_out/pdl_tests.txt
for i in 1 2; do
timeout 120 python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('off', d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])" >> gpurun_out/pdl_ab.txt
PDNN_PDL=1 timeout 120 python bench.py --steps 30 --warmup 5 2>gpurun_out/pdl_err.txt | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('on ', d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])" >> gpurun_out/pdl_ab.txt
done
cat gpurun_out/pdl_tests.txt gpurun_out/pdl_ab.txt; tail -3 gpurun_out/pdl_err.txt
