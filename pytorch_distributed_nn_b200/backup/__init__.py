"""Working counterparts of the reference's two parked experiments (pytorch_code/backup/): a `torch.distributed`
data-parallel MNIST-shape trainer (dist_nn_test.py) and an autograd hook walkthrough (back_ward_hook_test.py)."""
