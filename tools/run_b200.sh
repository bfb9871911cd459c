#!/bin/sh
# The reference's one committed launch line (pytorch_code/run_pytorch.sh: mpirun -n 3 ... --network=ResNet18
# --dataset=Cifar10 --batch-size=1024 --comm-type=Bcast --num-aggregate=5) on one 8xB200 box.
NGPU=${NGPU:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29500 \
    distributed_nn.py --network=ResNet18 --dataset=Cifar10 --batch-size=1024 --comm-type=Bcast --num-aggregate=5 \
    --colocate-worker --synthetic --max-steps ${STEPS:-100} "$@"
