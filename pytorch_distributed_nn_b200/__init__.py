"""pytorch_distributed_nn_b200 -- a B200-native synchronous-SGD parameter-server framework with the capabilities
of chao1224/pytorch_distributed_nn (master/worker roles, k-of-n aggregation with straggler kill, LeNet/ResNet/VGG,
torch state_dict checkpoints), built on hand-written sm_100a kernels (tcgen05/TMEM/TMA GEMM + implicit-GEMM conv,
fused layer kernels, NVLink peer-memory parameter-server kernels)."""
__version__ = "0.1.0"
