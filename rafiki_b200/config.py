"""Tunables of the serving path, same names as rafiki/config.py:14,17-18.

The reference sleeps 0.25 s between polls on both sides of its Redis hand-off and pops at
most 32 queries per loop (a structural ceiling of 128 queries/s per worker).  The in-process
hand-off here blocks on a condition variable instead of sleeping, so the sleeps survive only
as upper bounds on one wait, and the batch size is a per-worker parameter.
"""
import os

# Predictor
PREDICTOR_PREDICT_SLEEP = 0.25
# time after which Predictor.predict gives up on missing workers (the reference has none:
# rafiki/predictor/predictor.py:45 "TODO: add SLO") -- None keeps the reference's wait-forever
PREDICTOR_PREDICT_TIMEOUT = float(os.environ.get('PREDICTOR_PREDICT_TIMEOUT', 30.0))

# Inference worker
INFERENCE_WORKER_SLEEP = 0.25
INFERENCE_WORKER_PREDICT_BATCH_SIZE = int(os.environ.get('INFERENCE_WORKER_PREDICT_BATCH_SIZE', 65536))
