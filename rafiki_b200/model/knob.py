"""Knob types of the model plugin API (interface of rafiki/model/knob.py:4-198).

Knobs are hyper-parameter-search metadata: the serving path only needs model files that call
``get_knob_config()`` to import and to survive the JSON round trip the platform performs
(``serialize_knob_config`` / ``deserialize_knob_config``, rafiki/model/model.py:331-338).
"""
import abc
import json


class BaseKnob(abc.ABC):
    def __init__(self, knob_args=None):
        self._knob_args = dict(knob_args or {})

    def to_json(self):
        return json.dumps({'type': self.__class__.__name__, 'args': self._knob_args})

    @classmethod
    def from_json(cls, json_str):
        spec = json.loads(json_str)
        if 'type' not in spec or 'args' not in spec:
            raise ValueError('Invalid JSON representation of knob: {}.'.format(json_str))
        for clazz in (CategoricalKnob, IntegerKnob, FloatKnob, FixedKnob):
            if clazz.__name__ == spec['type']:
                return clazz(**spec['args'])
        raise ValueError('Invalid knob type: {}'.format(spec['type']))


def _value_type(value, allowed):
    for t in allowed:
        if isinstance(value, t) and not (t is int and isinstance(value, bool) and bool not in allowed):
            return t
    raise TypeError('Only the following types are supported: {}'.format(allowed))


class CategoricalKnob(BaseKnob):
    """One of ``values`` (all of one type among int, float, bool, str)."""

    def __init__(self, values):
        super().__init__({'values': values})
        if len(values) == 0:
            raise ValueError('Length of `values` should at least 1')
        first = bool if isinstance(values[0], bool) else _value_type(values[0], (int, float, str))
        if any(not isinstance(v, first) for v in values):
            raise TypeError('`values` should have elements of the same type')
        self._values = values
        self._value_type = first

    @property
    def value_type(self):
        return self._value_type

    @property
    def values(self):
        return self._values


class FixedKnob(BaseKnob):
    """A single fixed value (int, float, bool or str)."""

    def __init__(self, value):
        super().__init__({'value': value})
        self._value = value
        self._value_type = bool if isinstance(value, bool) else _value_type(value, (int, float, str))

    @property
    def value_type(self):
        return self._value_type

    @property
    def value(self):
        return self._value


class _RangeKnob(BaseKnob):
    _type = int

    def __init__(self, value_min, value_max, is_exp=False):
        super().__init__({'value_min': value_min, 'value_max': value_max, 'is_exp': is_exp})
        if value_min > value_max:
            raise ValueError('`value_max` should be at least `value_min`')
        self._value_min, self._value_max, self._is_exp = value_min, value_max, is_exp

    @property
    def value_min(self):
        return self._value_min

    @property
    def value_max(self):
        return self._value_max

    @property
    def is_exp(self):
        return self._is_exp


class IntegerKnob(_RangeKnob):
    """Any ``int`` in ``[value_min, value_max]``."""

    def __init__(self, value_min, value_max, is_exp=False):
        if not isinstance(value_min, int) or not isinstance(value_max, int):
            raise ValueError('`value_min` or `value_max` is not an `int`')
        super().__init__(value_min, value_max, is_exp)


class FloatKnob(_RangeKnob):
    """Any ``float`` in ``[value_min, value_max]``."""

    def __init__(self, value_min, value_max, is_exp=False):
        if not isinstance(value_min, (int, float)) or not isinstance(value_max, (int, float)):
            raise ValueError('`value_min` or `value_max` is not a number')
        super().__init__(float(value_min), float(value_max), is_exp)


def serialize_knob_config(knob_config):
    return json.dumps({name: knob.to_json() for (name, knob) in knob_config.items()})


def deserialize_knob_config(knob_config_str):
    return {name: BaseKnob.from_json(s) for (name, s) in json.loads(knob_config_str).items()}
