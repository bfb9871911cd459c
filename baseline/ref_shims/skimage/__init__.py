"""Import stub: the reference's MNIST reader imports skimage.util.random_noise (mnist/mnist.py:11) for an
augmentation helper that its training path never calls."""
