"""In-process replacement of the Redis hand-off (interface of rafiki/cache/cache.py:23-78).

Same seven methods, same semantics -- per-worker FIFO of queries, predictions looked up by
query id, a registry of the running workers of an inference job -- without the JSON text
round trips, and without two defects of the reference: ``pop_prediction_of_worker`` there
empties the whole list when it finds its prediction (``ltrim(key, i + 1, i)``, cache.py:74),
and the non-atomic ``lrange`` + ``ltrim`` pop (cache.py:47-50) races between replicas.  Here
both are single critical sections and a prediction is removed alone.

Waiting is on condition variables, so neither side sleeps 0.25 s between polls
(rafiki/config.py:14,17).
"""
import collections
import threading
import uuid


class Cache(object):
    def __init__(self, host=None, port=None):
        # host/port are accepted for signature compatibility; the hand-off is in-process
        self._lock = threading.Lock()
        self._queries_ready = threading.Condition(self._lock)
        self._predictions_ready = threading.Condition(self._lock)
        self._workers = collections.defaultdict(list)          # inference_job_id -> [worker_id] (insertion order)
        self._queries = collections.defaultdict(collections.deque)   # worker_id -> deque[(query_id, query)]
        self._predictions = collections.defaultdict(dict)      # worker_id -> {query_id: prediction}

    # ---- worker registry (cache.py:23-34)
    def add_worker_of_inference_job(self, worker_id, inference_job_id):
        with self._lock:
            if worker_id not in self._workers[inference_job_id]:
                self._workers[inference_job_id].append(worker_id)

    def delete_worker_of_inference_job(self, worker_id, inference_job_id):
        with self._lock:
            if worker_id in self._workers[inference_job_id]:
                self._workers[inference_job_id].remove(worker_id)

    def get_workers_of_inference_job(self, inference_job_id):
        with self._lock:
            return list(self._workers[inference_job_id])

    # ---- queries (cache.py:36-54)
    def add_query_of_worker(self, worker_id, query):
        query_id = str(uuid.uuid4())
        with self._lock:
            self._queries[worker_id].append((query_id, query))
            self._queries_ready.notify_all()
        return query_id

    def add_queries_of_worker(self, worker_id, queries):
        """Batch form of ``add_query_of_worker``: one critical section for the whole batch."""
        ids = [str(uuid.uuid4()) for _ in queries]
        with self._lock:
            self._queries[worker_id].extend(zip(ids, queries))
            self._queries_ready.notify_all()
        return ids

    def pop_queries_of_worker(self, worker_id, batch_size, timeout=None):
        """Up to ``batch_size`` oldest queries, FIFO.  With ``timeout`` it blocks until at least
        one query is there or the time is up (the reference returns at once and sleeps outside)."""
        with self._lock:
            q = self._queries[worker_id]
            if timeout and not q:
                self._queries_ready.wait_for(lambda: len(q) > 0, timeout)
            n = min(batch_size, len(q))
            items = [q.popleft() for _ in range(n)]
        return ([i for (i, _) in items], [x for (_, x) in items])

    # ---- predictions (cache.py:56-78)
    def add_prediction_of_worker(self, worker_id, query_id, prediction):
        with self._lock:
            self._predictions[worker_id][query_id] = prediction
            self._predictions_ready.notify_all()

    def add_predictions_of_worker(self, worker_id, query_ids, predictions):
        with self._lock:
            self._predictions[worker_id].update(zip(query_ids, predictions))
            self._predictions_ready.notify_all()

    def pop_prediction_of_worker(self, worker_id, query_id):
        """The prediction for ``query_id`` (removed), or ``None`` if it is not there yet."""
        with self._lock:
            return self._predictions[worker_id].pop(query_id, None)

    def wait_predictions_of_worker(self, worker_id, query_ids, timeout):
        """Block until every id has a prediction, then pop them all; ``None`` on timeout."""
        with self._lock:
            preds = self._predictions[worker_id]
            ok = self._predictions_ready.wait_for(lambda: all(i in preds for i in query_ids), timeout)
            if not ok:
                return None
            return [preds.pop(i) for i in query_ids]
