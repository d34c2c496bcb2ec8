"""No-op stand-in for the handful of TensorFlow calls the reference's driver SCRIPTS make before
they hand over to ``dino`` (dino_training.py:22-43: version check, disable_v2_behavior, set_seed).
It computes nothing.  Put ``dino_b200/compat`` on PYTHONPATH only when running those scripts."""
import sys as _sys

__version__ = "2.7.0"


def disable_v2_behavior():
    pass


class _Random:
    @staticmethod
    def set_seed(seed):
        from dino_b200.construction.neuralNetworks import set_seed as _s
        _s(seed)


random = _Random()


def set_random_seed(seed):
    _Random.set_seed(seed)


class _Compat:
    pass


compat = _Compat()
compat.v1 = _sys.modules[__name__]
_sys.modules[__name__ + ".compat"] = compat
_sys.modules[__name__ + ".compat.v1"] = _sys.modules[__name__]
