"""dino_b200: B200-native implementation of DINO's derivative-informed (H1) training step behind
the reference's ``dino.construction`` / ``dino.evaluation`` Python surface.  ``from dino_b200 import *``
gives what ``from dino import *`` gives in the reference (dino/__init__.py:18)."""
from .model import Adam, SGD, DinoModel, NetCore
from .construction import *
from . import evaluation
from . import inference
