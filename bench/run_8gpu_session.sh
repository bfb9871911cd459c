#!/bin/bash
# One 8-GPU lease: protocol checks, straggler study, link / comm sweeps, headline + reference benches (outputs in gpurun_out/r2/).
O=gpurun_out/r2; mkdir -p $O
F='Warning\|warn\|^\[W\|OMP_NUM\|\*\*\*\*\|^$'
run() { local n=$1 port=$2; shift 2; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@" 2>&1 | grep -v "$F"; }
run 8 29702 bench/ps_multigpu_check.py --network LeNet > $O/pscheck_lenet_8gpu.log
run 8 29703 bench/ps_multigpu_check.py --network ResNet18 --batch 32 > $O/pscheck_resnet18_8gpu.log
run 8 29704 bench/straggler_sweep.py configs/k_sweep/* configs/interval_sweep/* configs/time_cdf/* --master-only --out $O/straggler_study_8gpu_master_only.json > $O/straggler_study_8gpu.log
run 8 29701 bench/link_bench.py --mb 256 --out $O/link_8gpu.json > $O/link_8gpu.log
run 8 29705 bench.py --gpus 8 --steps 20 --warmup 5 --impl reference > $O/bench_ref_8gpu.json
run 8 29706 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_ours_8gpu.json
run 8 29707 bench.py --gpus 8 --steps 100 --warmup 5 --network VGG11 --master-only --num-aggregate 6 --inject-delay "exp:mean=2:p=0.3" > $O/bench_vgg11_k6of7_8gpu.json
run 8 29708 bench.py --gpus 8 --steps 100 --warmup 5 --network VGG11 --master-only --inject-delay "exp:mean=2:p=0.3" > $O/bench_vgg11_all7_8gpu.json
run 8 29709 bench.py --gpus 8 --steps 20 --warmup 5 --network ResNet50 > $O/bench_resnet50_8gpu.json
run 8 29710 bench.py --gpus 8 --steps 20 --warmup 5 --batch-size 128 > $O/bench_ours_b128_8gpu.json
run 8 29711 bench/comm_sweep.py --max-mb 1024 --iters 10 --out $O/comm_sweep_8gpu_nvls.json > $O/comm_sweep_8gpu_nvls.log
# 4 GPUs (0-3) and, next to them, 2 GPUs (4,5) for the sweeps
( export CUDA_VISIBLE_DEVICES=4,5; run 2 29721 bench/link_bench.py --mb 256 --out $O/link_2gpu.json > $O/link_2gpu.log; run 2 29722 bench/comm_sweep.py --max-mb 1024 --iters 10 --out $O/comm_sweep_2gpu_nvls.json > $O/comm_sweep_2gpu_nvls.log ) &
( export CUDA_VISIBLE_DEVICES=0,1,2,3; run 4 29731 bench/link_bench.py --mb 256 --out $O/link_4gpu.json > $O/link_4gpu.log; run 4 29732 bench/comm_sweep.py --max-mb 1024 --iters 10 --out $O/comm_sweep_4gpu_nvls.json > $O/comm_sweep_4gpu_nvls.log )
wait
tail -n 3 $O/pscheck_*.log | cut -c1-200; grep -c STUDY $O/straggler_study_8gpu.log; for f in $O/bench_*.json; do echo $f; tail -n 1 $f | cut -c1-330; done; cat $O/link_8gpu.log | tail -n 1 | cut -c1-700
