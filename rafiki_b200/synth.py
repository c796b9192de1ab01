"""Seeded synthetic queries and trials for tests and benchmarks.

There is no network, so no Fashion-MNIST and no trained checkpoints: queries are
synthetic ``uint8`` images of the reference's shapes (28x28 grey, 32x32x3 colour;
examples/datasets/image_classification/load_mnist_format.py:86-118) and trials are
random-initialised ``TfFeedForward`` networks whose knobs honour the reference's
ranges (examples/models/image_classification/TfFeedForward.py:20-28).

Batch-norm statistics are made consistent with the query distribution, as a trained
model's moving averages would be: "dead" features (constant border pixels) carry
``mean = 0, var = 0`` and the queries are zero there, so the ``rsqrt(eps)`` scale of
such features is exercised without producing absurd activations.

Seeds follow SURVEY.md 8(d): ``default_rng(seed)`` with seed 0 for data and
``1000 + t`` for trial ``t``.
"""
import numpy as np

NUM_CLASSES = 10          # rafiki/model/dataset.py:265 (classes of the image dataset)
BN_EPSILON = 1e-3


def dead_mask(d, frac=0.10, seed=7):
    """Boolean ``[d]``: features that are constant zero in the data."""
    rng = np.random.default_rng(seed)
    return rng.random(d) < frac


def make_queries(n, image_size=28, channels=1, kind='uniform', seed=0, dead=None):
    """``uint8 [n, S, S]`` (or ``[n, S, S, 3]``) query batch.

    kind='uniform': i.i.d. uniform 0..255 (throughput runs).
    kind='fashion': per-image random rectangle of N(128, 48^2) clipped to 0..255 on a
    zero background, which mimics the sparse borders of the reference's sample query
    (examples/scripts/quickstart.py:110-139).
    """
    rng = np.random.default_rng(seed)
    shape = (n, image_size, image_size) if channels == 1 else (n, image_size, image_size, channels)
    d = int(np.prod(shape[1:]))
    if kind == 'uniform':
        x = rng.integers(0, 256, size=shape, dtype=np.uint8)
    elif kind == 'fashion':
        x = np.zeros(shape, np.uint8)
        # rectangles may hang over the border, so that every pixel is covered reasonably often
        r0 = rng.integers(-(image_size // 2), image_size - 2, size=n)
        c0 = rng.integers(-(image_size // 2), image_size - 2, size=n)
        hh = rng.integers(image_size // 4, image_size + 1, size=n)
        ww = rng.integers(image_size // 4, image_size + 1, size=n)
        for i in range(n):
            ra, rb = max(r0[i], 0), min(r0[i] + hh[i], image_size)
            ca, cb = max(c0[i], 0), min(c0[i] + ww[i], image_size)
            if rb <= ra or cb <= ca:
                continue
            blk = (rb - ra, cb - ca) if channels == 1 else (rb - ra, cb - ca, channels)
            x[i, ra:rb, ca:cb] = np.clip(rng.normal(128.0, 48.0, size=blk), 0, 255).astype(np.uint8)
    else:
        raise ValueError('unknown query kind: {}'.format(kind))
    if dead is not None:
        x.reshape(n, d)[:, np.asarray(dead)] = 0
    return x


_FASHION_STATS = {}


def _fashion_stats(image_size, channels):
    """Per-feature mean / variance of the 'fashion' query distribution (2048-image sample)."""
    key = (image_size, channels)
    if key not in _FASHION_STATS:
        x = make_queries(2048, image_size, channels, 'fashion', seed=12345).reshape(2048, -1).astype(np.float64)
        _FASHION_STATS[key] = (x.mean(0), x.var(0))
    return _FASHION_STATS[key]


def make_trial(t, image_size=28, channels=1, hidden_layer_count=None, hidden_layer_units=None,
               num_classes=NUM_CLASSES, dead=None, data_kind='uniform'):
    """Parameter dict of trial ``t`` (the dict ``TfFeedForward.dump_parameters`` returns).

    Knobs not given are drawn from the reference's ranges: ``hidden_layer_count`` in
    {1, 2}, ``hidden_layer_units`` in 2..128 (TfFeedForward.py:23-24).  Dense kernels are
    Glorot-uniform ``[in, out]`` (Keras' default initialiser), biases N(0, 0.1^2),
    gamma ~ U(0.5, 1.5), beta ~ N(0, 0.1^2).
    """
    rng = np.random.default_rng(1000 + t)
    L = int(rng.integers(1, 3)) if hidden_layer_count is None else int(hidden_layer_count)
    H = int(rng.integers(2, 129)) if hidden_layer_units is None else int(hidden_layer_units)
    d = image_size * image_size * channels
    if dead is None:
        dead = np.zeros(d, bool)
    dead = np.asarray(dead)
    # moving statistics as training on this query distribution would leave them (+- jitter)
    if data_kind == 'uniform':
        mean = rng.normal(127.5, 6.0, size=d)
        var = rng.uniform(0.8, 1.2, size=d) * 5461.25
    else:
        m0, v0 = _fashion_stats(image_size, channels)
        mean = m0 * rng.uniform(0.95, 1.05, size=d)
        var = v0 * rng.uniform(0.8, 1.2, size=d)
    mean[dead] = 0.0
    var[dead] = 0.0
    dims = [d] + [H] * L + [num_classes]
    kernels, biases = [], []
    for fan_in, fan_out in zip(dims[:-1], dims[1:]):
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        kernels.append(rng.uniform(-lim, lim, size=(fan_in, fan_out)).astype(np.float32))
        biases.append(rng.normal(0.0, 0.1, size=fan_out).astype(np.float32))
    return {
        'image_size': int(image_size),
        'channels': int(channels),
        'hidden_layer_count': L,
        'hidden_layer_units': H,
        'bn_gamma': rng.uniform(0.5, 1.5, size=d).astype(np.float32),
        'bn_beta': rng.normal(0.0, 0.1, size=d).astype(np.float32),
        'bn_mean': mean.astype(np.float32),
        'bn_var': var.astype(np.float32),
        'bn_eps': BN_EPSILON,
        'kernels': kernels,
        'biases': biases,
    }


def trial_flops_per_query(params):
    """Algorithmic (unpadded) flops of one query through one trial:
    ``2 * sum(in * out)`` over the Dense layers."""
    return 2 * sum(int(k.shape[0]) * int(k.shape[1]) for k in params['kernels'])
