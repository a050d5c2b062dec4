"""B200-native hot path of derinworks/penr-oz-neural-network-torch (train step + batched inference).

``from penr_oz_neural_network_torch_b200 import NeuralNetworkModel`` is the drop-in for the
reference's ``from neural_net_model import NeuralNetworkModel``.
"""
from ._native import PozError, build, launch_count  # noqa: F401
from .engine import DeviceMLP, HostBatchFeeder, StackSpec, bind_host_to_gpu_numa_node  # noqa: F401
from .neural_net_model import (BatchNormLayer, EmbeddingLayer, FlattenLayer, Layer, LinearLayer,  # noqa: F401
                               MultiLayerPerceptron, NeuralNetworkModel, ReluLayer, SigmoidLayer, SoftmaxLayer,
                               TanhLayer)
