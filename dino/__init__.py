"""Alias package: ``from dino import *`` (what the reference's applications/ drivers do after
appending DINO_PATH) resolves to the B200-native implementation in ``dino_b200``."""
from dino_b200 import *            # noqa: F401,F403
from dino_b200 import construction, evaluation, inference  # noqa: F401
