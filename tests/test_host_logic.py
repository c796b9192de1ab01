"""Host-side mirror of the reference interface: cache hand-off, worker/predictor life cycle,
plugin API, knobs, resize.  CPU only (a fake model stands in for the CUDA trial)."""
import os
import pickle
import threading

import numpy as np
import pytest

from rafiki_b200.cache import Cache
from rafiki_b200.constants import TaskType
from rafiki_b200.db import Database
from rafiki_b200.model import (BaseModel, CategoricalKnob, FixedKnob, FloatKnob, IntegerKnob, dataset_utils,
                               deserialize_knob_config, load_model_class, serialize_knob_config)
from rafiki_b200.predictor import Predictor
from rafiki_b200.predictor.ensemble import ensemble_predictions
from rafiki_b200.worker import InferenceWorker

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

FAKE_MODEL = b'''
from rafiki_b200.model import BaseModel, FixedKnob
class Fake(BaseModel):
    @staticmethod
    def get_knob_config():
        return {'k': FixedKnob(1)}
    def __init__(self, **knobs):
        super().__init__(**knobs); self.p = None
    def train(self, uri): pass
    def evaluate(self, uri): return 1.0
    def predict(self, queries):
        if self.p.get('boom'): raise RuntimeError('boom')
        return [[q, self.p['tag']] for q in queries]
    def dump_parameters(self): return self.p
    def load_parameters(self, params): self.p = params
    def destroy(self): self.p = None
'''


def test_cache_fifo_and_prediction_lookup():
    c = Cache()
    c.add_worker_of_inference_job('w1', 'job')
    c.add_worker_of_inference_job('w2', 'job')
    assert c.get_workers_of_inference_job('job') == ['w1', 'w2']
    ids = [c.add_query_of_worker('w1', q) for q in 'abcde']
    got_ids, got = c.pop_queries_of_worker('w1', 3)
    assert got == list('abc') and got_ids == ids[:3]
    got_ids2, got2 = c.pop_queries_of_worker('w1', 32)
    assert got2 == list('de')
    assert c.pop_queries_of_worker('w1', 32) == ([], [])
    # predictions are removed one at a time (the reference's ltrim drops the whole list)
    for i, qid in enumerate(ids):
        c.add_prediction_of_worker('w1', qid, i)
    assert c.pop_prediction_of_worker('w1', ids[2]) == 2
    assert c.pop_prediction_of_worker('w1', ids[2]) is None
    assert c.pop_prediction_of_worker('w1', ids[0]) == 0
    assert c.pop_prediction_of_worker('w1', ids[4]) == 4
    assert c.pop_prediction_of_worker('w2', ids[1]) is None
    c.delete_worker_of_inference_job('w1', 'job')
    assert c.get_workers_of_inference_job('job') == ['w2']


def test_knob_config_round_trip():
    cfg = {'a': FixedKnob(2), 'b': IntegerKnob(1, 2), 'c': FloatKnob(1e-5, 1e-1, is_exp=True),
           'd': CategoricalKnob([16, 32, 64, 128])}
    back = deserialize_knob_config(serialize_knob_config(cfg))
    assert back['a'].value == 2 and back['b'].value_max == 2 and back['c'].is_exp and back['d'].values == [16, 32, 64, 128]
    with pytest.raises(ValueError):
        IntegerKnob(1.5, 2)


def test_resize_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, 'resize_reference.npz'))
    q = g['queries'].tolist()
    assert np.array_equal(dataset_utils.resize_as_images(q, [28, 28]), g['identity_28'])
    assert np.array_equal(dataset_utils.resize_as_images(q, [32, 32]), g['nearest_32'])
    assert np.array_equal(dataset_utils.resize_as_images(g['rgb_in'].tolist(), [32, 32]), g['rgb_identity_32'])
    assert dataset_utils.resize_as_images([[[256 + 7, -1], [0, 255]]], [2, 2]).tolist() == [[[7, 255], [0, 255]]]


def test_ensemble_edge_cases_without_gpu():
    assert ensemble_predictions([], TaskType.IMAGE_CLASSIFICATION) == []
    assert ensemble_predictions([[]], TaskType.IMAGE_CLASSIFICATION) == []
    pos = [[['a', 'b'], ['c']], [['x', 'y'], ['z']]]
    assert ensemble_predictions(pos, TaskType.POS_TAGGING) == [['a', 'b'], ['c']]
    assert ensemble_predictions([np.asarray([[1, 2]])], TaskType.POS_TAGGING) == [[1, 2]]


def _deploy(tmp_path, trial_params, task=TaskType.POS_TAGGING):
    db, cache = Database(), Cache()
    model = db.create_model(FAKE_MODEL, 'Fake', task)
    tj = db.create_train_job(task)
    stj = db.create_sub_train_job(tj.id, model.id)
    ij = db.create_inference_job(tj.id, 'predictor-svc')
    workers = []
    for i, p in enumerate(trial_params):
        path = tmp_path / 'trial{}.model'.format(i)
        path.write_bytes(pickle.dumps(p))
        trial = db.create_trial(stj.id, {'k': 1}, str(path))
        db.create_inference_job_worker('worker-{}'.format(i), ij.id, trial.id)
        workers.append(InferenceWorker('worker-{}'.format(i), cache=cache, db=db, batch_size=4))
    return db, cache, workers


def test_worker_predictor_round_trip(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    db, cache, workers = _deploy(tmp_path, [{'tag': 't0'}, {'tag': 't1'}])
    threads = [threading.Thread(target=w.start, daemon=True) for w in workers]
    for t in threads:
        t.start()
    pred = Predictor('predictor-svc', db=db, cache=cache)
    pred.start()
    for _ in range(200):
        if len(cache.get_workers_of_inference_job(pred._inference_job_id)) == 2:
            break
        threading.Event().wait(0.01)
    # non-image task: the ensemble returns the first trial's answer (ensemble.py:15-18)
    # (which worker is "first" follows registration order, as with the reference's Redis set)
    first = 't' + cache.get_workers_of_inference_job(pred._inference_job_id)[0][-1]
    assert pred.predict('q0') == {'prediction': ['q0', first]}
    out = pred.predict_batch(['a', 'b', 'c', 'd', 'e', 'f', 'g'])
    assert out == {'predictions': [[q, first] for q in 'abcdefg']}
    assert pred.predict_batch([]) == {'predictions': []}
    for w in workers:
        w.stop()
    for t in threads:
        t.join(2)
    assert workers[0].queries == 8 and workers[0].batches >= 3      # batch_size 4 -> 1 + 4 + 3
    assert cache.get_workers_of_inference_job(pred._inference_job_id) == []


def test_failed_predict_does_not_hang_the_predictor(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    db, cache, workers = _deploy(tmp_path, [{'tag': 't0', 'boom': True}])
    w = workers[0]
    with db:
        (job, trial) = w._read_worker_info()
        w._model = w._load_model(trial)
        cache.add_worker_of_inference_job('worker-0', job)
    pred = Predictor('predictor-svc', db=db, cache=cache)
    pred.start()
    t = threading.Thread(target=lambda: w.step(timeout=1.0), daemon=True)
    t.start()
    assert pred.predict('q') == {'prediction': None}
    t.join(2)


def test_failed_predict_on_an_image_task_answers_none(tmp_path, monkeypatch):
    """IMAGE_CLASSIFICATION: a worker whose predict() raised answers None; with no worker left the
    predictor returns {'prediction': None} instead of pushing None rows into the mean."""
    monkeypatch.chdir(tmp_path)
    db, cache, workers = _deploy(tmp_path, [{'tag': 't0', 'boom': True}], task=TaskType.IMAGE_CLASSIFICATION)
    w = workers[0]
    with db:
        (job, trial) = w._read_worker_info()
        w._model = w._load_model(trial)
        cache.add_worker_of_inference_job('worker-0', job)
    pred = Predictor('predictor-svc', db=db, cache=cache)
    pred.start()
    t = threading.Thread(target=lambda: (w.step(timeout=1.0), w.step(timeout=1.0)), daemon=True)
    t.start()
    assert pred.predict([[0] * 28] * 28) == {'prediction': None}
    assert pred.predict_batch([[[0] * 28] * 28] * 3) == {'predictions': [None, None, None]}
    t.join(3)


def test_binary_frames_ingress_without_gpu():
    from rafiki_b200.engine import frames_as_uint8
    a = np.arange(2 * 4 * 4, dtype=np.uint8).reshape(2, 4, 4)
    assert frames_as_uint8(a) is a or np.shares_memory(frames_as_uint8(a), a)
    b = frames_as_uint8(a.tobytes(), (4, 4))
    assert b.shape == (2, 4, 4) and np.array_equal(a, b)
    with pytest.raises(ValueError):
        frames_as_uint8(b'12345', (2, 2))
    with pytest.raises(ValueError):
        frames_as_uint8(b'1234')
    with pytest.raises(TypeError):
        frames_as_uint8(np.zeros((1, 4, 4), np.float32))


def test_resize_folds_into_the_first_layer():
    """Pillow-NEAREST resize + BatchNormalization + first Dense layer == one Dense layer on the raw frames
    (rafiki_b200.engine.fold_resize_into_params): checked against the oracle's resize-then-forward."""
    from oracle import feedforward as ff
    from rafiki_b200 import synth
    from rafiki_b200.engine import fold_resize_into_params
    for S, s_in, ch in ((32, 28, 1), (48, 28, 1), (32, 28, 3), (28, 40, 1), (32, 32, 1)):     # enlarge, reduce, same size
        p = synth.make_trial(3, S, ch, 2, 36, data_kind='fashion')
        q = synth.make_queries(40, s_in, ch, 'fashion', seed=2)
        want = ff.forward(p, ff.resize_as_images(q.tolist(), [S, S]), np.float64)
        pf = fold_resize_into_params(p, s_in, ch)
        assert pf['kernels'][0].shape[0] == s_in * s_in * ch and pf['image_size'] == s_in
        assert np.abs(ff.forward(pf, q, np.float64) - want).max() <= 1e-6


def test_plugin_contract_of_the_dropin_model_file():
    src = open(os.path.join(ROOT, 'examples', 'models', 'image_classification', 'TfFeedForward.py'), 'rb').read()
    cwd = os.getcwd()
    clazz = load_model_class(src, 'TfFeedForward')
    assert os.getcwd() == cwd and issubclass(clazz, BaseModel)
    cfg = clazz.get_knob_config()
    assert sorted(cfg) == ['batch_size', 'epochs', 'hidden_layer_count', 'hidden_layer_units', 'image_size',
                           'learning_rate']
    assert cfg['image_size'].value == 32 and cfg['hidden_layer_units'].value_max == 128
    deserialize_knob_config(serialize_knob_config(cfg))
    m = clazz(epochs=1, hidden_layer_count=1, hidden_layer_units=8, learning_rate=1e-3, batch_size=16, image_size=28)
    assert m.predict([]) == []
    from rafiki_b200.model import InvalidModelParamsException
    with pytest.raises(InvalidModelParamsException):
        m.load_parameters({'something': 'else'})
    with pytest.raises(InvalidModelParamsException):
        m.dump_parameters()
