"""The drop-in plugin, worker and predictor on the GPU, against the oracle's composition of the
reference path (oracle.feedforward.serve_batch / predictor_predict)."""
import os
import pickle
import threading

import numpy as np
import pytest
import torch

from oracle import feedforward as ff
from rafiki_b200 import engine, synth
from rafiki_b200.cache import Cache
from rafiki_b200.constants import TaskType
from rafiki_b200.db import Database
from rafiki_b200.model import serve_model_class
from rafiki_b200.predictor import Predictor
from rafiki_b200.worker import InferenceWorker

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODEL_FILE = os.path.join(ROOT, 'examples', 'models', 'image_classification', 'TfFeedForward.py')


def _knobs(p):
    return {'epochs': 1, 'hidden_layer_count': p['hidden_layer_count'], 'hidden_layer_units': p['hidden_layer_units'],
            'learning_rate': 1e-3, 'batch_size': 32, 'image_size': p['image_size']}


def _wire(p):
    keys = ('image_size', 'bn_gamma', 'bn_beta', 'bn_mean', 'bn_var', 'bn_eps', 'kernels', 'biases')
    return {k: p[k] for k in keys}


@pytest.mark.parametrize('S', [28, 32])
def test_dropin_model_file_through_the_plugin_harness(S, tmp_path, monkeypatch, golden_dir):
    monkeypatch.chdir(tmp_path)
    p = synth.make_trial(3, S, 1, 2, 36, data_kind='fashion')
    queries = synth.make_queries(40, 28, 1, 'fashion', seed=2).tolist()     # 28x28 queries, resized when S = 32
    queries[0] = np.load(os.path.join(golden_dir, 'resize_reference.npz'))['sample_query'].tolist()
    model, preds = serve_model_class(MODEL_FILE, 'TfFeedForward', TaskType.IMAGE_CLASSIFICATION, _wire(p), queries,
                                     _knobs(p))
    want = ff.serve_batch([p], queries)
    assert isinstance(preds, list) and len(preds) == 40 and len(preds[0]) == 10 and isinstance(preds[0][0], float)
    assert np.abs(np.asarray(preds) - np.asarray(want)).max() <= 2e-3
    again = pickle.loads(pickle.dumps(model.dump_parameters()))
    assert np.array_equal(again['kernels'][0], p['kernels'][0])
    model.destroy()


def test_default_knob_resize_runs_on_the_gpu_path(tmp_path, monkeypatch):
    """image_size 32 (the reference's FixedKnob, TfFeedForward.py:27) with 28 x 28 queries: tensor-core trials
    fold the resize into their first layer (no resize pass at all), the fp32 trial resizes with the CUDA
    kernel -- in neither case on the host."""
    from rafiki_b200 import _native
    monkeypatch.chdir(tmp_path)
    p = synth.make_trial(6, 32, 1, 2, 64, data_kind='fashion')
    frames = synth.make_queries(300, 28, 1, 'fashion', seed=9)                  # uint8 ndarray: binary ingress
    want = np.asarray(ff.serve_batch([p], frames.tolist()))
    r0, f0 = _native.launch_count('resize'), _native.launch_count('fused')
    model, preds = serve_model_class(MODEL_FILE, 'TfFeedForward', TaskType.IMAGE_CLASSIFICATION, _wire(p), frames, _knobs(p))
    assert np.abs(np.asarray(preds) - want).max() <= 2e-3
    assert _native.launch_count('resize') == r0 and _native.launch_count('fused') > f0
    assert np.abs(model.predict_array(frames) - want).max() <= 2e-3
    model.destroy()
    monkeypatch.setenv('RAFIKI_B200_COMPUTE', 'f32')
    model, preds = serve_model_class(MODEL_FILE, 'TfFeedForward', TaskType.IMAGE_CLASSIFICATION, _wire(p), frames, _knobs(p))
    assert np.abs(np.asarray(preds) - want).max() <= 1e-5
    assert _native.launch_count('resize_u8_staged') > r0
    model.destroy()
    # the pinned pipeline in front of an ensemble of such trials: frames in as they arrive, bytes or ndarray
    monkeypatch.delenv('RAFIKI_B200_COMPUTE')
    params = [synth.make_trial(t, 32, 1, 2, 48 + 16 * t, data_kind='fashion') for t in range(2)]
    ens = engine.Ensemble([engine.Trial(q) for q in params])
    pipe = engine.HostPipeline(ens, max_batch=512, depth=2, input_shape=(28, 28))
    got = pipe.serve(frames.tobytes())
    assert np.abs(got - np.asarray(ff.serve_batch(params, frames.tolist()))).max() <= 2e-3
    ens32 = engine.Ensemble([engine.Trial(q, compute='f32') for q in params])
    got32 = engine.HostPipeline(ens32, max_batch=512, depth=1, input_shape=(28, 28)).serve(frames)
    assert np.abs(got32 - np.asarray(ff.serve_batch(params, frames.tolist()))).max() <= 1e-5


def test_dropin_loads_a_reference_made_h5_parameter_dict(tmp_path, monkeypatch):
    """What a reference train worker stored: {'h5_model_base64': model.save() bytes} (TfFeedForward.py:102-118)."""
    import base64
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import h5_writer
    monkeypatch.chdir(tmp_path)
    p = synth.make_trial(5, 28, 1, 2, 64, data_kind='fashion')
    wire = {'h5_model_base64': base64.b64encode(h5_writer.keras_model_file(p, 28)).decode('utf-8')}
    queries = synth.make_queries(50, 28, 1, 'fashion', seed=4).tolist()
    model, preds = serve_model_class(MODEL_FILE, 'TfFeedForward', TaskType.IMAGE_CLASSIFICATION, wire, queries, _knobs(p))
    assert np.abs(np.asarray(preds) - np.asarray(ff.serve_batch([p], queries))).max() <= 2e-3
    assert np.array_equal(model.dump_parameters()['kernels'][1], p['kernels'][1])
    model.destroy()


def test_fp32_mode_of_the_dropin(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv('RAFIKI_B200_COMPUTE', 'f32')
    p = synth.make_trial(4, 28, 1, 1, 100)
    queries = synth.make_queries(33).tolist()
    model, preds = serve_model_class(MODEL_FILE, 'TfFeedForward', TaskType.IMAGE_CLASSIFICATION, _wire(p), queries,
                                     _knobs(p))
    assert np.abs(np.asarray(preds) - np.asarray(ff.serve_batch([p], queries))).max() <= 1e-5
    model.destroy()


def test_worker_predictor_ensemble_on_gpu(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    dead = synth.dead_mask(784)
    params = [synth.make_trial(t, 28, 1, L, H, dead=dead, data_kind='fashion')
              for t, (L, H) in enumerate([(2, 128), (1, 36), (2, 7)])]
    db, cache = Database(), Cache()
    task = TaskType.IMAGE_CLASSIFICATION
    model = db.create_model(open(MODEL_FILE, 'rb').read(), 'TfFeedForward', task)
    tj = db.create_train_job(task)
    stj = db.create_sub_train_job(tj.id, model.id)
    ij = db.create_inference_job(tj.id, 'predictor-svc')
    workers = []
    for i, p in enumerate(params):
        path = tmp_path / 'trial{}.model'.format(i)
        path.write_bytes(pickle.dumps(_wire(p)))
        trial = db.create_trial(stj.id, _knobs(p), str(path))
        db.create_inference_job_worker('worker-{}'.format(i), ij.id, trial.id)
        workers.append(InferenceWorker('worker-{}'.format(i), cache=cache, db=db, batch_size=64))
    threads = [threading.Thread(target=w.start, daemon=True) for w in workers]
    for t in threads:
        t.start()
    pred = Predictor('predictor-svc', db=db, cache=cache)
    pred.start()
    for _ in range(3000):
        if len(cache.get_workers_of_inference_job(ij.id)) == 3:
            break
        threading.Event().wait(0.01)
    queries = synth.make_queries(150, 28, 1, 'fashion', seed=5, dead=dead).tolist()
    one = pred.predict(queries[0])
    want_one = ff.predictor_predict(params, queries[0])
    assert sorted(one) == ['prediction'] and len(one['prediction']) == 10
    assert np.abs(np.asarray(one['prediction']) - np.asarray(want_one['prediction'])).max() <= 2e-3
    got = pred.predict_batch(queries)['predictions']
    want = ff.serve_batch(params, queries)
    assert np.abs(np.asarray(got) - np.asarray(want)).max() <= 2e-3
    # co-located trials: the predictor answers the batch with one fused launch
    trials = [engine.Trial(p) for p in params]
    pred.attach_ensemble(engine.Ensemble(trials), max_batch=128)
    fused = pred.predict_batch(queries)['predictions']
    assert np.abs(np.asarray(fused) - np.asarray(want)).max() <= 2e-3
    assert np.abs(np.asarray(fused) - np.asarray(got)).max() <= 1e-6
    for w in workers:
        w.stop()
    for t in threads:
        t.join(5)


def test_host_pipeline_overlapped_slots():
    dead = synth.dead_mask(784)
    params = [synth.make_trial(t, 28, 1, 2, 64, dead=dead) for t in range(3)]
    trials = [engine.Trial(p) for p in params]
    pipe = engine.HostPipeline(engine.Ensemble(trials), max_batch=512, depth=3)
    batches = [synth.make_queries(512 - 17 * i, 28, 1, 'uniform', seed=i, dead=dead) for i in range(3)]
    for i, b in enumerate(batches):
        pipe.stage(i, b)
        pipe.launch(i)
    for i, b in enumerate(batches):
        got = pipe.wait(i).numpy()
        want = ff.ensemble_mean_array([ff.forward(p, b, np.float32) for p in params])
        assert got.shape == (len(b), 10) and np.abs(got - want).max() <= 2e-3
