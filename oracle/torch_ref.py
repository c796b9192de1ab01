"""Second, independent restatement of the Keras forward pass, built from torch.nn
modules on the CPU.  TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

It exists because the Keras semantics in ``oracle/feedforward.py`` are restated
from the published layer definitions (TensorFlow 1.12 cannot be installed here):
two implementations that share no code must agree before either is trusted.
Follows ``examples/models/image_classification/TfFeedForward.py:141-157``.
"""
import numpy as np
import torch


def build_module(params, dtype=torch.float32):
    """Flatten -> BatchNorm1d(eval) -> [Linear + ReLU]*L -> Linear -> Softmax."""
    d = int(np.asarray(params['bn_gamma']).shape[0])
    bn = torch.nn.BatchNorm1d(d, eps=float(params.get('bn_eps', 1e-3)))
    with torch.no_grad():
        bn.weight.copy_(torch.as_tensor(np.asarray(params['bn_gamma'])))
        bn.bias.copy_(torch.as_tensor(np.asarray(params['bn_beta'])))
        bn.running_mean.copy_(torch.as_tensor(np.asarray(params['bn_mean'])))
        bn.running_var.copy_(torch.as_tensor(np.asarray(params['bn_var'])))
    layers = [torch.nn.Flatten(), bn]
    n_layers = len(params['kernels'])
    for i, (k, b) in enumerate(zip(params['kernels'], params['biases'])):
        k = np.asarray(k)
        lin = torch.nn.Linear(k.shape[0], k.shape[1])
        with torch.no_grad():
            lin.weight.copy_(torch.as_tensor(k).t())       # keras kernel is [in, out]
            lin.bias.copy_(torch.as_tensor(np.asarray(b)))
        layers.append(lin)
        layers.append(torch.nn.ReLU() if i < n_layers - 1 else torch.nn.Softmax(dim=-1))
    return torch.nn.Sequential(*layers).to(dtype).eval()


def forward(params, X, dtype=torch.float32):
    m = build_module(params, dtype)
    with torch.no_grad():
        x = torch.as_tensor(np.ascontiguousarray(np.asarray(X))).to(dtype)
        return m(x).numpy()
