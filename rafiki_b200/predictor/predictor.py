"""``Predictor`` with the reference's interface (rafiki/predictor/predictor.py:14-87).

``predict(query)`` keeps the reference's flow: hand the query to every running worker of the
inference job, gather one prediction per worker, ensemble, return ``{'prediction': ...}``.
``predict_batch(queries)``, an empty stub in the reference (``:85-87``), is implemented; when the
workers' trials share this process's GPU (``attach_ensemble``) a batch is answered by one fused
kernel launch instead of W hand-offs.  Besides the reference's nested lists it takes binary
batches -- a ``uint8`` ndarray ``[n, S, S(, ch)]`` or ``bytes`` of n frames -- which go from the
caller's memory into a pinned slot with one copy and no per-pixel Python work (the JSON text
hand-off of rafiki/cache/cache.py:36-45 is what this replaces); frames that are not the size the
trials were trained on are resized on the GPU.

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
        self._ensemble = ensemble
        self._max_batch = int(max_batch)
        self._pipelines = {}            # frame shape -> HostPipeline (pinned slots sized for that shape)
        self._pipeline = self._pipeline_for(self._ensemble_frame_shape())

    def _ensemble_frame_shape(self):
        e = self._ensemble
        return (e.image_size, e.image_size) + ((e.channels,) if e.channels > 1 else ())

    def _pipeline_for(self, frame_shape):
        from ..engine import HostPipeline
        frame_shape = tuple(int(v) for v in frame_shape)
        pipe = self._pipelines.get(frame_shape)
        if pipe is None:
            pipe = self._pipelines[frame_shape] = HostPipeline(self._ensemble, max_batch=self._max_batch, depth=2,
                                                               input_shape=frame_shape)
        return pipe

    def predict(self, query):
        predictions = self._fan_out([query])
        prediction = predictions[0] if len(predictions) > 0 else None
        return {'prediction': prediction}

    def predict_batch(self, queries, frame_shape=None, as_array=None):
        """Queries -> ``{'predictions': ...}``, one ensembled prediction per query.

        ``queries``: the reference's list of 2-D (or 3-D) pixel lists, or a binary batch: a ``uint8``
        ndarray ``[n, S, S(, ch)]``, or bytes of n frames with ``frame_shape`` = one frame's shape.
        ``as_array``: return a ``float32 [n, C]`` ndarray instead of nested lists (default: lists
        for list input, an array for binary input).
        """
        binary = isinstance(queries, (np.ndarray, bytes, bytearray, memoryview))
        if as_array is None:
            as_array = binary
        if len(queries) == 0:
            return {'predictions': np.zeros((0, 0), np.float32) if as_array else []}
        if self._pipeline is not None:
            from ..engine import frames_as_uint8
            from ..model.dataset import queries_to_uint8
            frames = frames_as_uint8(queries, frame_shape) if binary else queries_to_uint8(queries)
            pipe = self._pipeline_for(frames.shape[1:])
            out = []
            for lo in range(0, len(frames), pipe.max_batch):
                out.append(pipe.serve(frames[lo:lo + pipe.max_batch]))
            preds = out[0] if len(out) == 1 else np.concatenate(out)
            return {'predictions': preds if as_array else preds.tolist()}
        if isinstance(queries, (bytes, bytearray, memoryview)):
            from ..engine import frames_as_uint8
            queries = frames_as_uint8(queries, frame_shape)
        preds = self._fan_out(queries)
        return {'predictions': np.asarray(preds, dtype=np.float32) if as_array else preds}

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
            # a worker whose predict() raised answers the whole batch with None (worker/inference.py):
            # it is left out of the ensemble, as a worker that is not running would be
            if any(p is None for p in preds):
                logger.warning('worker %s failed on this batch; ensembling without it', worker_id)
                continue
            predictions_list.append(preds)
        if len(running_worker_ids) > 0 and len(predictions_list) == 0:
            return [None] * len(queries)       # every worker failed: {'prediction': None}, as for no workers
        return ensemble_predictions(predictions_list, self._task)

    def _read_predictor_info(self):
        inference_job = self._db.get_inference_job_by_predictor(self._service_id)
        train_job = self._db.get_train_job(inference_job.train_job_id)
        return (inference_job.id, train_job.task)
