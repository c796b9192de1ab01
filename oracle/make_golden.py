"""Generate the golden vectors under ``tests/golden/`` from the REFERENCE ITSELF.

TEST INFRASTRUCTURE ONLY.  Run in the build container, where ``/root/reference`` is
mounted (it does not exist on the GPU box, so only the vectors travel):

    python oracle/make_golden.py

What is pinned (outputs of the reference's own, unmodified functions):

* ``rafiki.predictor.ensemble.ensemble_predictions`` (rafiki/predictor/ensemble.py:6-34)
  on seeded [W][N][C] lists, its empty-input and non-image-task edge cases;
* ``ModelDatasetUtils.resize_as_images`` (rafiki/model/dataset.py:68-78) for the
  identity size (28 -> 28) as shipped, and for 28 -> 32 with the filter forced to
  NEAREST, which is what the pinned Pillow 5.3.0 default resolves to (this
  container's Pillow 12 defaults to BICUBIC -- SURVEY.md F6);
* the 28x28 sample query of examples/scripts/quickstart.py:110-139, extracted from
  the file's literal.

The Keras forward pass cannot be generated this way (tensorflow==1.12.0 is not
installable here); ``forward_restatement.npz`` holds the fp64 restatement's output on
fixed seeds as a regression vector only, and is labelled "unpinned" in its own keys.

The one harness shim: ``collections.Iterable`` was removed in Python 3.10 and
``rafiki/predictor/ensemble.py:2`` imports it, so the alias is restored before import.
"""
import ast
import collections
import collections.abc
import importlib.util
import os
import sys

import numpy as np

REF = '/root/reference'
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), 'tests', 'golden')


def _load_reference():
    collections.Iterable = collections.abc.Iterable          # harness shim (see docstring)
    sys.path.insert(0, REF)
    from rafiki.predictor.ensemble import ensemble_predictions   # noqa: E402
    from rafiki.constants import TaskType                         # noqa: E402
    # rafiki/model/__init__.py pulls in the advisor (needs baytune); load dataset.py alone
    spec = importlib.util.spec_from_file_location(
        'ref_dataset', os.path.join(REF, 'rafiki', 'model', 'dataset.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return ensemble_predictions, TaskType, mod.ModelDatasetUtils()


def _sample_query():
    """The literal ``queries`` list of examples/scripts/quickstart.py:110-139."""
    src = open(os.path.join(REF, 'examples', 'scripts', 'quickstart.py')).read()
    tree = ast.parse(src)
    for node in ast.walk(tree):
        if isinstance(node, ast.Assign) and getattr(node.targets[0], 'id', None) == 'queries':
            q = ast.literal_eval(node.value)
            if isinstance(q, list) and len(q) == 1 and len(q[0]) == 28:
                return q[0]
    raise RuntimeError('sample query not found')


def main():
    os.makedirs(OUT, exist_ok=True)
    ensemble_predictions, TaskType, dataset_utils = _load_reference()

    # ---- ensemble ------------------------------------------------------------------
    rng = np.random.default_rng(0)
    cases = {}
    for name, (w, n, c) in {'w1': (1, 5, 10), 'w2': (2, 33, 10), 'w3': (3, 257, 10),
                            'w8': (8, 64, 10), 'w5c3': (5, 17, 3)}.items():
        logits = rng.normal(size=(w, n, c)).astype(np.float32)
        p = np.exp(logits - logits.max(-1, keepdims=True))
        p = (p / p.sum(-1, keepdims=True)).astype(np.float32)
        # the workers hand the predictor Python lists of fp32 values (TfFeedForward.py:97)
        out = ensemble_predictions(p.tolist(), TaskType.IMAGE_CLASSIFICATION)
        cases[name + '_in'] = p
        cases[name + '_out'] = np.asarray(out, dtype=np.float64)
    # distinct one-hot rows: mean = counts / W
    onehot = np.zeros((4, 6, 10), np.float32)
    for wi in range(4):
        for ni in range(6):
            onehot[wi, ni, (wi * ni) % 10] = 1.0
    cases['onehot_in'] = onehot
    cases['onehot_out'] = np.asarray(
        ensemble_predictions(onehot.tolist(), TaskType.IMAGE_CLASSIFICATION), np.float64)
    # edge cases of ensemble.py:7-8 and :15-18
    assert ensemble_predictions([], TaskType.IMAGE_CLASSIFICATION) == []
    assert ensemble_predictions([[]], TaskType.IMAGE_CLASSIFICATION) == []
    pos = [[['a', 'b'], ['c']], [['x', 'y'], ['z']]]
    assert ensemble_predictions(pos, TaskType.POS_TAGGING) == [['a', 'b'], ['c']]
    cases['edge_empty_is_list'] = np.asarray([1])
    cases['edge_other_task_returns_first'] = np.asarray([1])
    np.savez_compressed(os.path.join(OUT, 'ensemble_reference.npz'), **cases)

    # ---- resize --------------------------------------------------------------------
    from PIL import Image
    q = _sample_query()
    imgs = rng.integers(0, 256, size=(6, 28, 28)).astype(np.uint8)
    queries = [q] + imgs.tolist()
    ident = dataset_utils.resize_as_images(queries, [28, 28])
    assert ident.dtype == np.uint8 and ident.shape == (7, 28, 28)
    # pinned-Pillow behaviour: same code path as dataset.py:76-78 with NEAREST explicit
    near = np.asarray([np.asarray(Image.fromarray(np.asarray(x, dtype=np.uint8))
                                  .resize([32, 32], Image.NEAREST)) for x in queries])
    # index table as PIL sees it, probed with a ramp image
    ramp = np.tile(np.arange(28, dtype=np.uint8), (28, 1))
    table = np.asarray(Image.fromarray(ramp).resize([32, 32], Image.NEAREST))[0]
    rgb = rng.integers(0, 256, size=(3, 32, 32, 3)).astype(np.uint8)
    rgb_out = dataset_utils.resize_as_images(rgb.tolist(), [32, 32])
    assert rgb_out.shape == (3, 32, 32, 3)
    np.savez_compressed(os.path.join(OUT, 'resize_reference.npz'),
                        sample_query=np.asarray(q, np.uint8), queries=np.asarray(queries, np.uint8),
                        identity_28=ident, nearest_32=near, nearest_table_28_32=table,
                        rgb_in=rgb, rgb_identity_32=rgb_out)

    # ---- forward restatement regression vector (UNPINNED: not from TensorFlow) -------
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import feedforward as ff
    from rafiki_b200 import synth
    dead = synth.dead_mask(784)
    fwd = {}
    for t, (L, H) in enumerate([(2, 128), (1, 36), (2, 7)]):
        params = synth.make_trial(t, 28, 1, L, H, dead=dead, data_kind='fashion')
        X = synth.make_queries(16, 28, 1, 'fashion', seed=t, dead=dead)
        X[0] = np.asarray(q, np.uint8)
        fwd['unpinned_t{}_probs_f64'.format(t)] = ff.forward(params, X, np.float64)
    np.savez_compressed(os.path.join(OUT, 'forward_restatement.npz'), **fwd)
    print('wrote', sorted(os.listdir(OUT)))


if __name__ == '__main__':
    main()
