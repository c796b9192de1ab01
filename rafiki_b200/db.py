"""In-memory stand-in for the seven ``Database`` reads the serving path performs
(rafiki/db/database.py; called from rafiki/worker/inference.py:78-105 and
rafiki/predictor/predictor.py:76-83).  The reference's PostgreSQL metadata store is out of
scope; a deployment passes its own ``db`` object with these methods instead.
"""
import types
import uuid


class Database(object):
    def __init__(self):
        self._rows = {k: {} for k in ('model', 'train_job', 'sub_train_job', 'trial', 'inference_job', 'worker')}
        self._predictor_of_job = {}

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    @staticmethod
    def _row(**fields):
        fields.setdefault('id', str(uuid.uuid4()))
        return types.SimpleNamespace(**fields)

    # ---- writes used by the deployment harness
    def create_model(self, model_file_bytes, model_class, task):
        r = self._row(model_file_bytes=model_file_bytes, model_class=model_class, task=task)
        self._rows['model'][r.id] = r
        return r

    def create_train_job(self, task):
        r = self._row(task=task)
        self._rows['train_job'][r.id] = r
        return r

    def create_sub_train_job(self, train_job_id, model_id):
        r = self._row(train_job_id=train_job_id, model_id=model_id)
        self._rows['sub_train_job'][r.id] = r
        return r

    def create_trial(self, sub_train_job_id, knobs, params_file_path, score=0.0):
        r = self._row(sub_train_job_id=sub_train_job_id, knobs=knobs, params_file_path=params_file_path, score=score)
        self._rows['trial'][r.id] = r
        return r

    def create_inference_job(self, train_job_id, predictor_service_id):
        r = self._row(train_job_id=train_job_id, predictor_service_id=predictor_service_id)
        self._rows['inference_job'][r.id] = r
        self._predictor_of_job[predictor_service_id] = r.id
        return r

    def create_inference_job_worker(self, service_id, inference_job_id, trial_id):
        r = self._row(service_id=service_id, inference_job_id=inference_job_id, trial_id=trial_id)
        self._rows['worker'][service_id] = r
        return r

    # ---- reads of the serving path
    def get_inference_job_by_predictor(self, predictor_service_id):
        return self._rows['inference_job'].get(self._predictor_of_job.get(predictor_service_id))

    def get_train_job(self, id):
        return self._rows['train_job'].get(id)

    def get_inference_job_worker(self, service_id):
        return self._rows['worker'].get(service_id)

    def get_inference_job(self, id):
        return self._rows['inference_job'].get(id)

    def get_trial(self, id):
        return self._rows['trial'].get(id)

    def get_sub_train_job(self, id):
        return self._rows['sub_train_job'].get(id)

    def get_model(self, id):
        return self._rows['model'].get(id)
