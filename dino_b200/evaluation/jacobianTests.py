"""Jacobian error metrics of compute_jacobian_errors (jacobianTests.py:37-130), fed from arrays.

The reference draws prior samples and evaluates the true Jacobian with a hIPPYlib PDE model; that
side needs FEniCS and is out of scope.  Here the truth is passed in as arrays and the prediction
comes from the CUDA Jacobian chain.  For DinoModel / InferenceOracle predictions every O(d_Q^2 d_M) contraction of the
seven metrics also runs on the device (``jacobian_error_metrics_device``); ``jacobian_error_metrics`` is the host
restatement of the reference's lines for arbitrary callables."""
import math

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


def _grams_device(X, Y):
    """G[b] = X[b] Y[b]^T for device tensors [n, q, m]: one contraction of the CUDA library per sample
    (dino_gemm_nt; the tcgen05 fp32x3 engine when the row pitch allows TMA, else the fp32 FFMA engine)."""
    import ctypes as C
    import torch
    from .. import _lib
    from ..model import _stream
    lib = _lib.load()
    n, q, m = X.shape
    G = torch.empty(n, q, (q + 3) & ~3, dtype=torch.float32, device=X.device)
    eng = _lib.PREC_FP32X3 if m % 4 == 0 else _lib.PREC_FP32
    st = _stream()
    for b in range(n):
        _lib.check(lib.dino_gemm_nt(eng, C.c_void_p(X[b].data_ptr()), m, C.c_void_p(Y[b].data_ptr()), m,
                                    C.c_void_p(G[b].data_ptr()), G.shape[2], q, q, m, st), "dino_gemm_nt")
    return G[:, :, :q]


def jacobian_error_metrics_device(J_pred, J_true):
    """The seven metrics of jacobianTests.py:104-126 for a batch, with every O(d_Q^2 d_M) contraction on the device.

    With D = J_pred - J_true and the d_Q x d_Q Gram matrices DD^T, E = D J^T, B = J J^T (three device contractions per
    sample; P P^T = B + E + E^T + D D^T), everything the reference forms in d_M x d_M or through an SVD of the
    [d_Q, d_M] truth reduces to d_Q x d_Q algebra (host, float64): B = U S^2 U^T gives U and the singular values,
    V = J^T U S^-1, so D V = E U S^-1, J V = U S, and
      J^T J error     ||D^T P + J^T D||_F^2 = <DD^T, PP^T> + <B, DD^T> + 2 tr(E F),  F = D P^T = E + DD^T
      J J^T error     ||E + E^T + DD^T||_F / ||B||_F        (= the reduced one: U is square)
      reduced J       ||D V||_F / ||J||_F ;  nullspace  sqrt(||D||_F^2 - ||D V||_F^2) / ||J||_F   (J (I - V V^T) = 0)
      reduced J^T J   ||(DV)^T (PV) + (JV)^T (DV)||_F / ||S^2||_F
    Written in differences D, none of these subtracts two nearly equal large numbers.  Needs d_Q <= d_M and a truth of
    full row rank (as the reference's PDE Jacobians are); returns {metric: ndarray[n]}."""
    import torch
    P, J = J_pred, J_true
    n, q, m = J.shape
    if q > m:
        raise ValueError("jacobian_error_metrics_device needs d_Q <= d_M")
    D = (P - J).contiguous()
    DD = _grams_device(D, D).double().cpu().numpy()
    E = _grams_device(D, J).double().cpu().numpy()
    B = _grams_device(J, J).double().cpu().numpy()
    out = {k: np.empty(n) for k in ('Jerror', 'JTJerror', 'JJTerror', 'rJTJerror', 'rJJTerror', 'reducedJerror', 'nullspace_error')}
    fro = np.linalg.norm
    for b in range(n):
        dd, e, bb = 0.5 * (DD[b] + DD[b].T), E[b], 0.5 * (B[b] + B[b].T)
        pp = bb + e + e.T + dd
        s2, U = np.linalg.eigh(bb)
        s = np.sqrt(np.maximum(s2, 0.0))
        nJ, nB = math.sqrt(np.trace(bb)), fro(bb)
        F = e + dd                                   # D P^T
        DV = e @ U / s                               # D V
        JV = U * s                                   # J V = U S
        PV = JV + DV
        out['Jerror'][b] = math.sqrt(max(np.trace(dd), 0.0)) / nJ
        out['JTJerror'][b] = math.sqrt(max((dd * pp).sum() + (bb * dd).sum() + 2.0 * np.trace(e @ F), 0.0)) / nB
        out['JJTerror'][b] = fro(e + e.T + dd) / nB
        out['rJJTerror'][b] = out['JJTerror'][b]
        out['reducedJerror'][b] = fro(DV) / nJ
        out['nullspace_error'][b] = math.sqrt(max(np.trace(dd) - (DV * DV).sum(), 0.0)) / nJ
        out['rJTJerror'][b] = fro(DV.T @ PV + JV.T @ DV) / fro(s2)
    return out


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
        model = getattr(oracle, '_jac', oracle)           # InferenceOracle wraps a DinoModel with a Jacobian output
        J_true = np.asarray(truth['J_data'][:n_samples])
        if hasattr(model, 'predict_device') and J_true.shape[1] <= J_true.shape[2]:
            # prediction from the CUDA chain, the O(d_Q^2 d_M) contractions of every metric on the device
            import torch
            P_dev = model.predict_device(m)[1]
            J_dev = torch.from_numpy(np.ascontiguousarray(J_true, dtype=np.float32)).to(P_dev.device)
            errs = jacobian_error_metrics_device(P_dev, J_dev)
            for k in keys:
                master_logger[file][k] = [float(v) for v in errs[k]]
            continue
        if hasattr(oracle, 'J'):
            J_pred = np.stack([oracle.J(m[i]) for i in range(n_samples)])
        else:
            J_pred = oracle.predict(m)[1]
        for i in range(n_samples):
            errs = jacobian_error_metrics(np.asarray(J_pred[i], dtype=np.float64), np.asarray(truth['J_data'][i]))
            for k in keys:
                master_logger[file][k].append(errs[k])
    return master_logger
