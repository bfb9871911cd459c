# PDL on/off where launch edges weigh most: batch 128 ResNet-18, batch 1024 ResNet-50, VGG-11 (1 GPU)
for cfg in "ResNet18 128" "ResNet50 1024" "VGG11 1024"; do
set -- $cfg
for pdl in 0 1; do
PDNN_PDL=$pdl timeout 150 python bench.py --network $1 --batch-size $2 --steps 30 --warmup 5 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1 b$2 pdl=$pdl', d['ms_per_step'], d['value'])" >> gpurun_out/pdl_ab_small.txt
done
done
cat gpurun_out/pdl_ab_small.txt
