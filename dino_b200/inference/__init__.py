"""Array methods of the reference's InferenceOracle (inferenceOracle.py:105-139)."""
from .inferenceOracle import InferenceOracle, inferenceOracleSettings
