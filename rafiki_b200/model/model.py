"""The model plugin contract, kept verbatim from the reference (rafiki/model/model.py:20-127):
a model file defines a ``BaseModel`` subclass; the platform instantiates it with the trial's
knobs, calls ``load_parameters`` with the unpickled parameter dict and then ``predict`` on
batches of queries.
"""
import abc
import json
import os
import pickle
import sys
import uuid
from importlib import import_module

from ..constants import TaskType
from .knob import BaseKnob, deserialize_knob_config, serialize_knob_config


class InvalidModelClassException(Exception): pass
class InvalidModelParamsException(Exception): pass


class BaseModel(abc.ABC):
    """Rafiki's base model class.  Subclasses call ``super().__init__(**knobs)`` first."""

    def __init__(self, **knobs):
        pass

    @staticmethod
    def get_knob_config():
        raise NotImplementedError()

    @abc.abstractmethod
    def train(self, dataset_uri):
        raise NotImplementedError()

    @abc.abstractmethod
    def evaluate(self, dataset_uri):
        raise NotImplementedError()

    @abc.abstractmethod
    def predict(self, queries):
        raise NotImplementedError()

    @abc.abstractmethod
    def dump_parameters(self):
        raise NotImplementedError()

    @abc.abstractmethod
    def load_parameters(self, params):
        raise NotImplementedError()

    @abc.abstractmethod
    def destroy(self):
        pass


def load_model_class(model_file_bytes, model_class, temp_mod_name=None):
    """rafiki/model/model.py:221-242: write the model file's bytes to ``<name>.py`` in the
    working directory, import it, pull the class out, delete the file."""
    if temp_mod_name is None:
        temp_mod_name = 'm' + uuid.uuid4().hex
    temp_model_file_name = '{}.py'.format(temp_mod_name)
    with open(temp_model_file_name, 'wb') as f:
        f.write(model_file_bytes)
    cwd = os.getcwd()
    added = cwd not in sys.path
    if added:
        sys.path.insert(0, cwd)
    try:
        mod = import_module(temp_mod_name)
        clazz = getattr(mod, model_class)
    finally:
        os.remove(temp_model_file_name)
        if added:
            sys.path.remove(cwd)
    return clazz


def serve_model_class(model_file_path, model_class, task, parameters, queries, knobs):
    """The inference half of ``test_model_class`` (rafiki/model/model.py:183-210): pickle round
    trip of the parameters, fresh instance, ``load_parameters``, ``predict``, JSON check, and the
    single-trial ensemble the reference runs at the end."""
    from ..predictor.ensemble import ensemble_predictions
    try:
        with open(model_file_path, 'rb') as f:
            py_model_class = load_model_class(f.read(), model_class)
        if not issubclass(py_model_class, BaseModel):
            raise Exception('Model should extend `rafiki.model.BaseModel`')
        knob_config = py_model_class.get_knob_config()
        for (name, knob) in knob_config.items():
            if not isinstance(knob, BaseKnob):
                raise Exception('Invalid knob {}'.format(name))
        deserialize_knob_config(serialize_knob_config(knob_config))
        parameters = pickle.loads(pickle.dumps(parameters))
        model_inst = py_model_class(**knobs)
        model_inst.load_parameters(parameters)
        predictions = model_inst.predict(queries)
        for prediction in predictions:
            json.dumps(prediction)
        predictions = ensemble_predictions([predictions], task)
        return model_inst, predictions
    except Exception as e:
        raise InvalidModelClassException(e)


__all__ = ['BaseModel', 'InvalidModelClassException', 'InvalidModelParamsException', 'load_model_class',
           'serve_model_class', 'TaskType']
