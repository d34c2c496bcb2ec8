"""The reference's gradient tests compare against the PDE adjoint gradient and need FEniCS +
hIPPYlib (gradientTests.py:24-126); the PDE side is out of scope here."""


def gradient_error_test(*args, **kwargs):
    raise NotImplementedError('gradient_error_test needs the PDE model (FEniCS/hIPPYlib), which this build does not include')


def gradient_timing_test(*args, **kwargs):
    raise NotImplementedError('gradient_timing_test is an unfinished stub in the reference (gradientTests.py:104-126)')
