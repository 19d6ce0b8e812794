"""fedb200 - a Blackwell (B200, sm_100a) native federated-learning framework.

Capabilities follow amolahinge/739-839-federated-learning-using-grpc:
the ``federated.Trainer`` gRPC contract, a primary/backup coordinator pair, workers that
train the CIFAR model zoo locally, synchronous FedAvg, the ``-c`` and ``-a`` flags and the
``{'net','acc','epoch'}`` checkpoint layout.  The data plane is hand-written CUDA.
"""
__version__ = "0.1.0"
