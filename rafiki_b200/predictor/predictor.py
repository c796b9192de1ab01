"""``Predictor`` with the reference's interface (rafiki/predictor/predictor.py:14-87).

``predict(query)`` keeps the reference's flow: hand the query to every running worker of the
inference job, gather one prediction per worker, ensemble, return ``{'prediction': ...}``.
``predict_batch(queries)``, an empty stub in the reference (``:85-87``), is implemented; when the
workers' trials share this process's GPU (``attach_ensemble``) a batch is answered by one fused
kernel launch instead of W hand-offs.

Differences on purpose: waiting is event-driven with a timeout (the reference polls every
0.25 s forever, ``:45-59``), and a worker that fails to answer raises instead of hanging.
"""
import logging

import numpy as np

from ..cache import Cache
from ..config import PREDICTOR_PREDICT_TIMEOUT
from .ensemble import ensemble_predictions

logger = logging.getLogger(__name__)


class PredictorTimeoutError(RuntimeError): pass


class Predictor(object):
    def __init__(self, service_id, db=None, cache=None):
        if db is None:
            raise ValueError('Predictor needs a db object (rafiki_b200.db.Database or the platform\'s)')
        if cache is None:
            cache = Cache()
        self._service_id = service_id
        self._db = db
        self._cache = cache
        self._ensemble = None
        self._pipeline = None

    def start(self):
        with self._db:
            (self._inference_job_id, self._task) = self._read_predictor_info()

    def attach_ensemble(self, ensemble, max_batch=65536):
        """Serve batches through a co-located :class:`rafiki_b200.engine.Ensemble`."""
        from ..engine import HostPipeline
        self._ensemble = ensemble
        self._pipeline = HostPipeline(ensemble, max_batch=max_batch, depth=2)

    def predict(self, query):
        predictions = self._fan_out([query])
        prediction = predictions[0] if len(predictions) > 0 else None
        return {'prediction': prediction}

    def predict_batch(self, queries):
        """List of queries -> ``{'predictions': [...]}``, one ensembled prediction per query."""
        if len(queries) == 0:
            return {'predictions': []}
        if self._pipeline is not None:
            from ..model.dataset import dataset_utils
            s = self._ensemble_image_size()
            x = dataset_utils.resize_as_images(queries, [s, s] if s else None)
            out = []
            for lo in range(0, len(x), self._pipeline.max_batch):
                out.append(self._pipeline.serve(x[lo:lo + self._pipeline.max_batch]))
            return {'predictions': np.concatenate(out).tolist()}
        return {'predictions': self._fan_out(queries)}

    def _ensemble_image_size(self):
        d = self._ensemble.d
        for ch in (1, 3):
            s = int(round((d / ch) ** 0.5))
            if s * s * ch == d:
                return s
        return None

    def _fan_out(self, queries):
        running_worker_ids = self._cache.get_workers_of_inference_job(self._inference_job_id)
        worker_to_query_ids = {}
        for worker_id in running_worker_ids:
            worker_to_query_ids[worker_id] = self._cache.add_queries_of_worker(worker_id, queries)
        predictions_list = []
        for worker_id in running_worker_ids:
            preds = self._cache.wait_predictions_of_worker(worker_id, worker_to_query_ids[worker_id],
                                                           PREDICTOR_PREDICT_TIMEOUT)
            if preds is None:
                raise PredictorTimeoutError('worker {} did not answer within {} s'.format(
                    worker_id, PREDICTOR_PREDICT_TIMEOUT))
            predictions_list.append(preds)
        return ensemble_predictions(predictions_list, self._task)

    def _read_predictor_info(self):
        inference_job = self._db.get_inference_job_by_predictor(self._service_id)
        train_job = self._db.get_train_job(inference_job.train_job_id)
        return (inference_job.id, train_job.task)
