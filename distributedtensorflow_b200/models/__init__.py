"""Model families of the reference: the MNIST CNN (run_mnist_distributed.py:46-70), the MNIST MLP
(templates/00_mnist_replica.py:138-164) and a wide MLP for communication-bound benchmarking."""
