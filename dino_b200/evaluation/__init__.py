"""Same exports as the reference's dino/evaluation/__init__.py:19-23."""
from .gradientTests import gradient_error_test, gradient_timing_test
from .jacobianTests import compute_jacobian_errors, jacobian_error_metrics
from .surrogatePostProcessing import evaluateJacobianNetwork
