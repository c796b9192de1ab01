"""NumPy restatement of the reference serving path.  TEST INFRASTRUCTURE ONLY.

Every function names the reference lines it follows.  Citations are relative to
``/root/reference``.  Parity status: see ``oracle/__init__.py`` -- the resize and
ensemble restatements are pinned to outputs of the reference's own functions
(``tests/golden/``); the Keras forward is "parity unpinned" (TensorFlow 1.12 is a
third-party dependency that cannot be installed here).

A trial is described by a plain dict (the same dict the drop-in model pickles):

    {'image_size': S, 'bn_gamma','bn_beta','bn_mean','bn_var': float32 [D],
     'bn_eps': 1e-3, 'kernels': [float32 [in,out], ...], 'biases': [float32 [out], ...]}

``kernels[-1]``/``biases[-1]`` are the softmax layer; the others are ReLU layers.
"""
import numpy as np

IMAGE_CLASSIFICATION = 'IMAGE_CLASSIFICATION'   # rafiki/constants.py:59-60
KERAS_PREDICT_BATCH = 32                        # keras Model.predict default batch_size (TfFeedForward.py:95)
BN_EPSILON = 1e-3                               # keras BatchNormalization default (TfFeedForward.py:148)


# --------------------------------------------------------------------------- preprocess
def nearest_index_table(n_in, n_out):
    """Source index per destination index of Pillow's NEAREST resize
    (default filter of the pinned Pillow 5.3.0, rafiki/model/dataset.py:77):
    ``src = floor((dst + 0.5) * n_in / n_out)``.  Verified against PIL by
    ``oracle/make_golden.py`` (tests/golden/resize_nearest.npz)."""
    dst = np.arange(n_out, dtype=np.float64)
    src = np.floor((dst + 0.5) * (float(n_in) / float(n_out))).astype(np.int64)
    return np.minimum(src, n_in - 1)


def resize_as_images(images, image_size):
    """rafiki/model/dataset.py:68-78.

    ``np.asarray(x, dtype=np.uint8)`` per image (values outside 0..255 wrap, as the
    pinned NumPy 1.14.5 cast does), PIL ``resize(image_size)`` with the pinned default
    filter NEAREST, stack to ``uint8 [N, S, S]`` (``[N, S, S, 3]`` for RGB lists).
    """
    out = []
    for x in images:
        a = np.asarray(x)
        if a.dtype.kind == 'f':
            a = a.astype(np.int64)
        a = (a.astype(np.int64) & 0xFF).astype(np.uint8)
        # PIL size is (width, height): image_size[0] indexes columns
        rows = nearest_index_table(a.shape[0], int(image_size[1]))
        cols = nearest_index_table(a.shape[1], int(image_size[0]))
        out.append(a[rows][:, cols])
    return np.asarray(out)


# --------------------------------------------------------------------------- forward
def bn_affine(params, dtype=np.float32):
    """Inference-time BatchNormalization as a per-feature affine
    (keras.layers.BatchNormalization defaults, TfFeedForward.py:148):
    ``inv = gamma * rsqrt(var + eps)``, ``y = x * inv + (beta - mean * inv)``."""
    g = np.asarray(params['bn_gamma'], dtype)
    b = np.asarray(params['bn_beta'], dtype)
    m = np.asarray(params['bn_mean'], dtype)
    v = np.asarray(params['bn_var'], dtype)
    eps = dtype(params.get('bn_eps', BN_EPSILON))
    inv = (g / np.sqrt(v + eps)).astype(dtype)
    return inv, (b - m * inv).astype(dtype)


def softmax(z):
    """tf.nn.softmax along the class axis (TfFeedForward.py:154-157)."""
    z = z - z.max(axis=-1, keepdims=True)
    e = np.exp(z)
    return e / e.sum(axis=-1, keepdims=True)


def forward(params, X, dtype=np.float32, batch_size=None):
    """The Keras graph of ``_build_model`` (TfFeedForward.py:141-157) run the way
    ``self._model.predict(X)`` runs it (``:95``): uint8 -> float cast, Flatten
    (row-major, channels last), BatchNormalization (inference), ``relu(x @ K + b)``
    per hidden layer, ``softmax(x @ K_out + b_out)``.

    ``batch_size=32`` reproduces Keras' internal batching; rows are independent so
    the result does not depend on it.  ``dtype=np.float64`` gives the high-precision
    variant used to bound the fp32 oracle's own rounding.
    """
    X = np.asarray(X)
    n = X.shape[0]
    x = X.reshape(n, -1)
    inv, shift = bn_affine(params, dtype)
    kernels = [np.asarray(k, dtype) for k in params['kernels']]
    biases = [np.asarray(b, dtype) for b in params['biases']]
    c = kernels[-1].shape[1]
    out = np.empty((n, c), dtype)
    step = n if not batch_size else int(batch_size)
    for s in range(0, n, max(step, 1)):
        h = x[s:s + step].astype(dtype)
        h = h * inv + shift
        for k, b in zip(kernels[:-1], biases[:-1]):
            h = np.maximum(h @ k + b, dtype(0))
        out[s:s + step] = softmax(h @ kernels[-1] + biases[-1])
    return out


def predict(params, queries, batch_size=KERAS_PREDICT_BATCH):
    """``TfFeedForward.predict`` (TfFeedForward.py:89-97): list of queries in,
    ``probs.tolist()`` out."""
    s = int(params['image_size'])
    X = resize_as_images(queries, [s, s])
    if len(queries) == 0:
        return []
    return forward(params, X, np.float32, batch_size).tolist()


# --------------------------------------------------------------------------- ensemble
def ensemble_predictions(predictions_list, task):
    """rafiki/predictor/ensemble.py:6-34, restated.

    Empty input -> ``[]``; IMAGE_CLASSIFICATION -> per-query mean over trials of the
    class-probability vectors, computed in float64 exactly as ``np.mean`` over the
    transposed list does; any other task -> the first trial's predictions.  ndarray
    results are turned back into lists.
    """
    if len(predictions_list) == 0 or len(predictions_list[0]) == 0:
        return []
    if task == IMAGE_CLASSIFICATION:
        stacked = np.transpose(np.asarray(predictions_list), axes=[1, 0, 2])   # [N, W, C]
        predictions = [np.mean(p, axis=0) for p in stacked]
    else:
        predictions = predictions_list[0]
    if isinstance(predictions, np.ndarray):
        predictions = predictions.tolist()
    predictions = list(predictions) if not isinstance(predictions, list) else predictions
    for i in range(len(predictions)):
        if isinstance(predictions[i], np.ndarray):
            predictions[i] = predictions[i].tolist()
    return predictions


def ensemble_mean_array(probs):
    """Array form of the IMAGE_CLASSIFICATION branch above: ``probs`` is
    ``[W, N, C]``; returns float64 ``[N, C]`` (ensemble.py:13-14)."""
    p = np.asarray(probs, dtype=np.float64)
    return np.stack([np.mean(q, axis=0) for q in np.transpose(p, (1, 0, 2))]) if p.shape[1] else \
        np.zeros((0, p.shape[2]))


def predictor_predict(trial_params, query, task=IMAGE_CLASSIFICATION):
    """``Predictor.predict`` (rafiki/predictor/predictor.py:31-74) with the cache
    round trip removed: every worker answers the same single query, answers are
    wrapped as ``[[pred_w]]`` (``:64-67``), ensembled, and element 0 returned."""
    predictions_list = [[predict(p, [query])[0]] for p in trial_params]
    predictions = ensemble_predictions(predictions_list, task)
    return {'prediction': predictions[0] if len(predictions) > 0 else None}


def serve_batch(trial_params, queries, task=IMAGE_CLASSIFICATION):
    """The whole hot path on a batch, composed as ``test_model_class`` composes it
    (rafiki/model/model.py:200-210): ``predict`` per trial, then
    ``ensemble_predictions`` over the trials' lists."""
    return ensemble_predictions([predict(p, queries) for p in trial_params], task)
