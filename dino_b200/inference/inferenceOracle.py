"""InferenceOracle without the PDE model: ``J``, ``Jvp``, ``Jtvp`` of a trained network
(inferenceOracle.py:105-139) served by the CUDA Jacobian chain.  The reference builds three extra
TF graphs (Jt-product, full Jacobian, J-product by contraction with the full Jacobian, :79-98);
here ``J`` is ``dino_jacobian``, ``Jtvp`` a reverse chain with the single seed row qhat (s = 1) and ``Jvp`` one
forward-mode sweep (``dino_jvp``): neither product materialises the Jacobian.  The PDE-backed methods (cost/misfit/gradient with
'PDE', hIPPYlib vectors) need FEniCS and are not provided."""
import numpy as np

from ..construction.jacobianConstruction import equip_model_with_full_jacobian


def inferenceOracleSettings(settings={}):
    """Keys and defaults of inferenceOracle.py:26-43."""
    settings['misfit_method'] = 'NN'
    settings['cost_method'] = 'NN'
    settings['gradient_method'] = 'NN'
    settings['use_regularization'] = True
    settings['riesz_operator'] = 'regularization'
    return settings


class InferenceOracle:
    def __init__(self, hpModelWrapper, neural_operator, settings=inferenceOracleSettings()):
        assert neural_operator is not None
        self.modelwrapper = hpModelWrapper          # may be None: only the array methods are served
        self.neural_operator = neural_operator
        self.settings = settings
        self.observable_shape = tuple(neural_operator.outputs[0].shape[1:])
        self.dQ = int(np.prod(self.observable_shape))
        self.parameter_shape = tuple(neural_operator.inputs[0].shape[1:])
        self.dM = int(np.prod(self.parameter_shape))
        self._jac = neural_operator if len(neural_operator.outputs) == 2 else \
            equip_model_with_full_jacobian(neural_operator)
        self._jt = None                             # the same network with a one-row output seed (Jtvp)

    @staticmethod
    def _rows(a):
        a = np.asarray(a)
        return a[None, :] if a.ndim == 1 else a

    def J(self, mi):
        """Full Jacobian(s): [dQ, dM] for one sample, [n, dQ, dM] for several (inferenceOracle.py:127-139)."""
        mi = self._rows(mi)
        J = self._jac.predict(mi)[1]
        return J[0] if mi.shape[0] == 1 else J

    def Jvp(self, mi, mhat, sess=None):
        """J(m) mhat (inferenceOracle.py:116-125).  The reference contracts the full batch Jacobian with mhat
        (:96-98, "fwd mode was difficult to get to work in tf"); here it is one forward-mode sweep (dino_jvp)."""
        mi, mhat = self._rows(mi), self._rows(mhat)
        return self._jac.jvp(mi, mhat)[0]

    def Jtvp(self, mi, qhat):
        """J(m)^T qhat (inferenceOracle.py:79-82,105-114): a reverse chain seeded with the single row qhat."""
        mi, qhat = self._rows(mi), self._rows(qhat)
        if self._jt is None:
            from ..model import DinoModel
            self._jt = DinoModel(self._jac.core, 'out_reduced', seed_dim=1, output_basis=np.zeros((self.dQ, 1), np.float32),
                                 precision=self._jac.precision)
        self._jt.output_basis = np.ascontiguousarray(qhat[0].astype(np.float32)[:, None])        # Phi = qhat as a column
        return self._jt.predict(mi)[1][0, 0]

    def _pde_only(self, *a, **k):
        raise NotImplementedError('needs the hIPPYlib/FEniCS model wrapper, which this build does not include')

    cost = misfit = variational_gradient = gradient = _pde_only
