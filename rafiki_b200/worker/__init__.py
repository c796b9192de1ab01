from .inference import InferenceWorker, InvalidWorkerException
