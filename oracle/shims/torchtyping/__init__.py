"""Oracle-only stand-in for `torchtyping` (annotations only in the reference:
learned_dict.py:6, fista.py:8).  TEST INFRASTRUCTURE."""


class _TensorTypeMeta(type):
    def __getitem__(cls, item):
        return cls


class TensorType(metaclass=_TensorTypeMeta):
    pass
