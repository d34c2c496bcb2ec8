"""Jacobian error metrics of compute_jacobian_errors (jacobianTests.py:37-130), fed from arrays.

The reference draws prior samples and evaluates the true Jacobian with a hIPPYlib PDE model; that
side needs FEniCS and is out of scope.  Here the truth is passed in as arrays and the prediction
comes from the CUDA Jacobian chain (``oracle.J``-style callables or DinoModel.predict)."""
import numpy as np


def reduce_dict(master_logger, reduction='mean'):
    reduced_logger = {}
    for file in master_logger.keys():
        reduced_logger[file] = {}
        for key in master_logger[file].keys():
            if reduction == 'mean':
                reduced_logger[file][key] = np.mean(master_logger[file][key])
    return reduced_logger


def jacobian_error_metrics(J_full_pred, J_full):
    """The seven per-sample relative errors of jacobianTests.py:104-126."""
    norm = np.linalg.norm
    U, _, VT = np.linalg.svd(J_full, full_matrices=False)
    JtJ, JJt = J_full.T @ J_full, J_full @ J_full.T
    PtP, PPt = J_full_pred.T @ J_full_pred, J_full_pred @ J_full_pred.T
    return {
        'Jerror': norm(J_full_pred - J_full) / norm(J_full),
        'JTJerror': norm(PtP - JtJ) / norm(JtJ),
        'JJTerror': norm(PPt - JJt) / norm(JJt),
        'reducedJerror': norm(U.T @ J_full_pred @ VT.T - U.T @ J_full @ VT.T) / norm(U.T @ J_full @ VT.T),
        'nullspace_error': norm(J_full_pred - J_full_pred @ VT.T @ VT) / norm(J_full),
        'rJTJerror': norm(VT @ PtP @ VT.T - VT @ JtJ @ VT.T) / norm(VT @ JtJ @ VT.T),
        'rJJTerror': norm(U.T @ PPt @ U - U.T @ JJt @ U) / norm(U.T @ JJt @ U),
    }


def compute_jacobian_errors(truth, oracle_dictionary, n_samples=100, rank=50, p=5):
    """``truth``: dict with 'm_data' [n, d_M] and 'J_data' [n, d_Q, d_M] (replaces the PDE model
    wrapper of the reference).  ``oracle_dictionary``: {name: object with .J(m) or DinoModel with a
    full-Jacobian output}.  Returns {name: {metric: [per-sample values]}} like the reference."""
    keys = ['Jerror', 'JTJerror', 'JJTerror', 'rJTJerror', 'rJJTerror', 'reducedJerror', 'nullspace_error']
    n_samples = min(n_samples, truth['m_data'].shape[0])
    m = np.asarray(truth['m_data'][:n_samples])
    master_logger = {}
    for file, oracle in oracle_dictionary.items():
        master_logger[file] = {k: [] for k in keys}
        if hasattr(oracle, 'J'):
            J_pred = np.stack([oracle.J(m[i]) for i in range(n_samples)])
        else:
            J_pred = oracle.predict(m)[1]
        for i in range(n_samples):
            errs = jacobian_error_metrics(np.asarray(J_pred[i], dtype=np.float64), np.asarray(truth['J_data'][i]))
            for k in keys:
                master_logger[file][k].append(errs[k])
    return master_logger
