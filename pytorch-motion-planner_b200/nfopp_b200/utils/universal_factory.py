"""Name -> class reflection factory with the reference's semantics (utils/universal_factory.py:13-44):
a dict carrying "name"/"type" is built recursively, extra keyword arguments the constructor does not accept
are silently dropped, and an unknown class name raises KeyError("Unknown class ...")."""
from inspect import signature

from .attribute_dict import AttributeDict

PARAMETER_ATTRIBUTE_NAMES = ["parameters", "params", "param", "parameter"]
CLASS_NAME_ATTRIBUTES = ["name", "type"]


class UniversalFactory(object):
    def __init__(self, classes):
        self._classes = {cls.__name__: cls for cls in classes}

    def make_from_parameters(self, parameters, **kwargs):
        if not isinstance(parameters, (AttributeDict, dict)):
            return parameters
        ctor = None
        for key in CLASS_NAME_ATTRIBUTES:
            if key in parameters:
                if parameters[key] not in self._classes:
                    raise KeyError(f"Unknown class {parameters[key]}")
                ctor = self._classes[parameters[key]]
        if ctor is None:
            return parameters
        merged = dict(kwargs)
        merged.update({k: self.make_from_parameters(v) for k, v in parameters.items()})
        accepted = signature(ctor).parameters
        for key in PARAMETER_ATTRIBUTE_NAMES:
            if key in accepted:
                merged[key] = parameters
        return ctor(**{k: v for k, v in merged.items() if k in accepted})

    @staticmethod
    def kwargs_function(function):
        accepted = signature(function).parameters

        def call(*args, **kwargs):
            return function(*args, **{k: v for k, v in kwargs.items() if k in accepted})
        return call
