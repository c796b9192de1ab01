"""``InferenceWorker`` with the reference's interface (rafiki/worker/inference.py:19-105).

Same life cycle -- read the worker's trial, register with the inference job, load the model
class and its pickled parameters, then loop: pop a batch of queries, ``model.predict``, push the
predictions.  The loop blocks on the hand-off instead of sleeping 0.25 s per iteration
(``:65``), takes batches up to ``INFERENCE_WORKER_PREDICT_BATCH_SIZE`` (32 in the reference,
rafiki/config.py:18) and does not write whole batches to the log (``:48-49,59-60``).  A failed
``predict`` is logged and answered with ``None`` predictions so the predictor is not left
waiting (the reference drops the batch, ``:52-56``, and the predictor hangs).
"""
import logging
import pickle
import threading
import traceback

from ..cache import Cache
from ..config import INFERENCE_WORKER_PREDICT_BATCH_SIZE, INFERENCE_WORKER_SLEEP
from ..model import load_model_class

logger = logging.getLogger(__name__)


class InvalidWorkerException(Exception): pass


class InferenceWorker(object):
    def __init__(self, service_id, cache=None, db=None, batch_size=None):
        if db is None:
            raise ValueError('InferenceWorker needs a db object')
        if cache is None:
            cache = Cache()
        self._cache = cache
        self._db = db
        self._service_id = service_id
        self._model = None
        self._batch_size = batch_size or INFERENCE_WORKER_PREDICT_BATCH_SIZE
        self._stop = threading.Event()
        self.batches = 0
        self.queries = 0

    def start(self):
        logger.info('Starting inference worker for service of id {}...'.format(self._service_id))
        with self._db:
            (inference_job_id, trial_id) = self._read_worker_info()
            self._model = self._load_model(trial_id)
            # register only once the model can answer, so a predictor never fans out to a cold worker
            self._cache.add_worker_of_inference_job(self._service_id, inference_job_id)
        while not self._stop.is_set():
            self.step(timeout=INFERENCE_WORKER_SLEEP)

    def step(self, timeout=None):
        """One iteration of the serving loop; returns the number of queries answered."""
        (query_ids, queries) = self._cache.pop_queries_of_worker(self._service_id, self._batch_size, timeout)
        if len(queries) == 0:
            return 0
        predictions = None
        try:
            predictions = self._model.predict(queries)
        except Exception:
            logger.error('Error while making predictions:')
            logger.error(traceback.format_exc())
        if predictions is None:
            predictions = [None] * len(queries)
        self._cache.add_predictions_of_worker(self._service_id, query_ids, predictions)
        self.batches += 1
        self.queries += len(queries)
        return len(queries)

    def stop(self):
        self._stop.set()
        with self._db:
            (inference_job_id, _) = self._read_worker_info()
        self._cache.delete_worker_of_inference_job(self._service_id, inference_job_id)
        if self._model is not None:
            self._model.destroy()
            self._model = None

    def _load_model(self, trial_id):
        trial = self._db.get_trial(trial_id)
        sub_train_job = self._db.get_sub_train_job(trial.sub_train_job_id)
        model = self._db.get_model(sub_train_job.model_id)
        clazz = load_model_class(model.model_file_bytes, model.model_class)
        model_inst = clazz(**trial.knobs)
        with open(trial.params_file_path, 'rb') as f:
            parameters = pickle.loads(f.read())
        model_inst.load_parameters(parameters)
        return model_inst

    def _read_worker_info(self):
        worker = self._db.get_inference_job_worker(self._service_id)
        if worker is None:
            raise InvalidWorkerException()
        inference_job = self._db.get_inference_job(worker.inference_job_id)
        return (inference_job.id, worker.trial_id)
