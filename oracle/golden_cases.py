"""Golden-vector case list shared by oracle/make_golden.py and the tests -- TEST INFRASTRUCTURE.

STEP_CASES follow the Linear+activation rows of the reference's own ``test_train`` table
(test_neural_net_model.py:161-179) and ``test_compute_output`` table (:133-142), plus an
MNIST-shaped stack (BASELINE.json configs[1]) at a reduced batch.
"""

def _c(name, sizes, algos, optimizer, bias, batch, salt, steps=3, lr=0.01, decay=0.9, l2=0.001,
       **kw):
    d = dict(name=name, sizes=sizes, algos=algos, optimizer=optimizer, bias=bias, batch=batch,
             salt=salt, steps=steps, lr=lr, decay=decay, l2=l2)
    d.update(kw)
    return d


STEP_CASES = [
    _c("sig_sig_sgd",      [9, 9, 9],    ["sigmoid"] * 2,                None,   "zeros",  7, 1),
    _c("relu_softmax_adam", [9, 9, 9],   ["relu", "softmax"],            "adam", "zeros",  8, 2),
    _c("tanh_tanh_sgd",    [9, 9, 9],    ["tanh"] * 2,                   None,   "random", 5, 3),
    _c("relu_sig_adam",    [18, 9, 3],   ["relu", "sigmoid"],            "adam", "random", 6, 4),
    _c("sig_lin_softmax",  [9, 18, 9],   ["sigmoid", "linear", "softmax"], None, "zeros",  9, 5),
    _c("sig_sig_wide_out", [4, 8, 16],   ["sigmoid"] * 2,                None,   "zeros",  4, 6),
    _c("relu_relu_softmax", [3, 3, 3, 3], ["relu", "relu", "softmax"],   "adam", "random", 11, 7),
    _c("relu_relu_sgd",    [18, 9, 3],   ["relu"] * 2,                   None,   "zeros",  3, 8),
    _c("relu_tanh_adam",   [9, 18, 9],   ["relu", "tanh"],               "adam", "zeros",  10, 9),
    _c("softmax_only",     [9, 9],       ["softmax"],                    "adam", "zeros",  2, 10),
    _c("tanh_softmax",     [4, 8, 16],   ["tanh", "softmax"],            None,   "random", 13, 11),
    _c("default_relu",     [3, 3],       None,                           "adam", "zeros",  1, 12),
    _c("nobias_relu_softmax", [12, 7, 5], ["relu", "softmax"],           "adam", "",       6, 13),
    _c("lin_lin_tail",     [6, 5, 4, 3], ["relu", "linear", "linear"],   None,   "random", 5, 14),
    _c("odd_sizes_adam",   [37, 65, 33, 17], ["tanh", "sigmoid", "softmax"], "adam", "random", 67, 15,
       steps=4),
    _c("dropout_relu_softmax", [9, 12, 9], ["relu", "softmax"], "adam", "random", 16, 16,
       dropout=0.3),
    _c("dropout_tanh_mse", [8, 16, 8, 4], ["tanh", "sigmoid", "relu"],   None,   "zeros",  12, 17,
       dropout=0.25),
    _c("mnist_b64_adam",   [784, 256, 128, 10], ["relu", "relu", "softmax"], "adam", "zeros", 64, 18,
       steps=2, subsample=61),
    _c("mnist_b32_sgd",    [784, 256, 128, 10], ["relu", "relu", "softmax"], None, "random", 32, 19,
       steps=2, subsample=61),
]

TRAIN_API_CASES = [
    dict(name="relu_softmax_adam", sizes=[9, 9, 9], algos=["relu", "softmax"], optimizer="adam",
         bias="zeros", salt=21, seed=5, epochs=3, batch=16, lr=0.01, decay=0.9, l2=0.001),
    dict(name="sig_sig_sgd", sizes=[9, 9, 9], algos=["sigmoid"] * 2, optimizer=None,
         bias="random", salt=22, seed=6, epochs=2, batch=None, lr=0.05, decay=0.8, l2=0.0, extra=3),
    dict(name="relu_tanh_adam", sizes=[9, 18, 9], algos=["relu", "tanh"], optimizer="adam",
         bias="zeros", salt=23, seed=7, epochs=120, batch=8, lr=0.01, decay=0.99, l2=0.001),
]

INIT_CASES = [
    dict(name="xavier_zeros", sizes=[9, 9, 9], algos=["relu", "softmax"], optimizer="adam",
         weight_algo="xavier", bias="zeros", confidence=1.0, seed=0),
    dict(name="he_random_conf", sizes=[18, 9, 3], algos=["relu", "tanh"], optimizer=None,
         weight_algo="he", bias="random", confidence=0.5, seed=1),
    dict(name="gaussian_mixed", sizes=[9, 18, 9], algos=["sigmoid", "linear", "softmax"],
         optimizer="adam", weight_algo="gaussian", bias="random", confidence=0.7, seed=2),
    dict(name="he_nobias", sizes=[5, 6, 7, 8], algos=["tanh", "relu", "sigmoid"],
         optimizer="adam", weight_algo="he", bias="", confidence=1.0, seed=3),
    dict(name="emb_bn_he", sizes=[27, 10, 30, 64, 27], algos=["embedding", "linear", "batchnorm", "tanh", "linear", "softmax"],
         optimizer="adam", weight_algo="he", bias="random", confidence=0.1, seed=4),
    dict(name="emb_hier_xavier", sizes=[9, 2, 4, 9, 18, 9],
         algos=["embedding", "linear", "batchnorm", "tanh", "flatten", "linear", "softmax"],
         optimizer=None, weight_algo="xavier", bias="zeros", confidence=1.0, seed=5),
]


# Stacks with embedding / flatten / batchnorm entries: the remaining rows of
# the reference's test_train table (test_neural_net_model.py:180-185), the model of
# test_torch_backward.py:121-135, and batchnorm-only MLPs.  `tokens` = indices per sample.
def _n(name, sizes, algos, optimizer, bias, batch, salt, tokens=0, steps=3, lr=0.01, decay=0.9, l2=0.001, **kw):
    d = dict(name=name, sizes=sizes, algos=algos, optimizer=optimizer, bias=bias, batch=batch, salt=salt,
             tokens=tokens, steps=steps, lr=lr, decay=decay, l2=l2)
    d.update(kw)
    return d


NET_CASES = [
    _n("bn_tanh_softmax", [12, 16, 7], ["linear", "batchnorm", "tanh", "softmax"], "adam", "random", 9, 31),
    _n("bn_relu_bn_sig_mse", [10, 14, 9, 5], ["linear", "batchnorm", "relu", "linear", "batchnorm", "sigmoid", "linear"],
       None, "zeros", 11, 32),
    _n("emb_tanh_softmax", [9, 2, 6, 18, 9], ["embedding", "tanh", "softmax"], "adam", "zeros", 8, 33, tokens=3),
    _n("emb_lin_bn_tanh_softmax", [9, 2, 6, 18, 9], ["embedding", "linear", "batchnorm", "tanh", "softmax"], None,
       "random", 10, 34, tokens=3),
    _n("emb_hier_flatten", [9, 2, 4, 9, 18, 9],
       ["embedding", "linear", "batchnorm", "tanh", "flatten", "linear", "softmax"], None, "zeros", 6, 35, tokens=4),
    _n("makemore_bn", [27, 10, 30, 64, 27], ["embedding", "linear", "batchnorm", "tanh", "linear", "softmax"], "adam",
       "random", 32, 36, tokens=3, steps=2),
    _n("emb_dropout", [9, 4, 12, 10, 9], ["embedding", "linear", "batchnorm", "relu", "softmax"], None, "random", 12,
       37, tokens=3, dropout=0.25),
]


# Dropout in distribution (neural_net_model.py:393-395): the reference's cost and mean |weight gradient| of one
# step over many F.dropout seeds.  The device draws its masks from Philox, so single steps cannot agree; the
# means over the same number of seeds must, within the standard error.
DROPSTAT_CASES = [
    dict(name="relu_relu_softmax_p20", sizes=[64, 48, 32, 10], algos=["relu", "relu", "softmax"], optimizer=None,
         bias="random", batch=512, salt=41, dropout=0.2, seeds=64),
]
