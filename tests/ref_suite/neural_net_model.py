"""Shim that stands where the reference's ``neural_net_model.py`` stands on ``sys.path``: the reference's own
test files (and its ``main.py``) do ``from neural_net_model import ...`` and get the B200 drop-in instead.
Used by tests/test_gpu_ref_suite.py only."""
from penr_oz_neural_network_torch_b200.neural_net_model import (  # noqa: F401
    BatchNormLayer, EmbeddingLayer, FlattenLayer, Layer, LinearLayer, MultiLayerPerceptron, NeuralNetworkModel,
    ReluLayer, SigmoidLayer, SoftmaxLayer, TanhLayer)

__poz_dropin__ = True
