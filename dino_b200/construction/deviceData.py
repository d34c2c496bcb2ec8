"""Device-side versions of the data contractions that precede training.

The reference builds its targets in host numpy float64: ``remap_jacobian_data`` (dataUtilities.py:492-510) loops over
the samples forming J_i = U_i diag(sigma_i) V_i^T, ``prune_the_data`` (jacobianTrainingDriver.py:93-185) restricts
m, q, J to the reduced bases with einsums.  At the full-Jacobian configuration that is 13.8 GB of output.  Here each
contraction is a call of this library's GEMM entry points on device tensors (dino_gemm_nt / dino_gemm_tn_acc: the
tcgen05 fp32x3 engine when the row pitches allow TMA, the fp32 FFMA engine otherwise); host code only marshals.
Results are float32 - what training consumes (Keras casts the reference's float64 arrays on feed) - and agree with the
reference's float64 outputs to ~1e-6 relative (tests/test_gpu_data.py checks 1e-5 against the golden vectors).
The host versions in dataUtilities.py / jacobianTrainingDriver.py stay the bit-exact float64 mirrors.
"""
import ctypes as C

import numpy as np

from .. import _lib


def _dev(a):
    import torch
    from ..model import _device
    if isinstance(a, torch.Tensor):
        return a.to(device=_device(), dtype=torch.float32).contiguous()
    return torch.from_numpy(np.ascontiguousarray(np.asarray(a), dtype=np.float32)).to(_device())


def _engine(*pitches):
    return _lib.PREC_FP32X3 if all(p % 4 == 0 for p in pitches) else _lib.PREC_FP32


def gemm_nt(A, Bt, out=None):
    """C[M, N] = A[M, K] Bt[N, K]^T on the device (2-D contiguous fp32 tensors)."""
    import torch
    from ..model import _stream
    lib = _lib.load()
    M, K = A.shape
    N = Bt.shape[0]
    assert Bt.shape[1] == K
    ldc = (N + 3) & ~3
    Cm = torch.empty(M, ldc, dtype=torch.float32, device=A.device) if out is None else out
    _lib.check(lib.dino_gemm_nt(_engine(K, ldc), C.c_void_p(A.data_ptr()), K, C.c_void_p(Bt.data_ptr()), K,
                                C.c_void_p(Cm.data_ptr()), Cm.shape[1], M, N, K, _stream()), "dino_gemm_nt")
    return Cm[:, :N]


def gemm_tn(X, Y):
    """C[n_in, n_out] = X[rows, n_in]^T Y[rows, n_out] on the device."""
    import torch
    from ..model import _stream
    lib = _lib.load()
    rows, n_in = X.shape
    n_out = Y.shape[1]
    Cm = torch.zeros(n_in, n_out, dtype=torch.float32, device=X.device)
    _lib.check(lib.dino_gemm_tn_acc(_engine(n_in, n_out), C.c_void_p(X.data_ptr()), n_in, C.c_void_p(Y.data_ptr()), n_out,
                                    C.c_void_p(Cm.data_ptr()), rows, n_in, n_out, _stream()), "dino_gemm_tn_acc")
    return Cm


def remap_jacobian_data_device(data_dict, as_numpy=True):
    """J_i = U_i diag(sigma_i) V_i^T for every sample (dataUtilities.py:492-510): one device contraction per sample,
    [d_Q, k] x [d_M, k]^T.  Returns the dict of remap_jacobian_data with J_data float32 (numpy, or a device tensor)."""
    import torch
    U, sg, V = _dev(data_dict['U_data']), _dev(data_dict['sigma_data']), _dev(data_dict['V_data'])
    n, dQ, k = U.shape
    dM = V.shape[1]
    Us = (U * sg[:, None, :]).contiguous()                  # column scaling (elementwise)
    ld = (dM + 3) & ~3
    J = torch.empty(n, dQ, ld, dtype=torch.float32, device=U.device)
    for i in range(n):
        gemm_nt(Us[i], V[i], out=J[i])
    J = J[:, :, :dM]
    J = J.contiguous() if ld != dM else J
    return {'m_data': data_dict['m_data'], 'q_data': data_dict['q_data'], 'J_data': J.cpu().numpy() if as_numpy else J}


def prune_the_data_device(projector_dict, train_dict, test_dict, input_pruning=True, output_pruning=True, as_numpy=True):
    """prune_the_data (jacobianTrainingDriver.py:93-185) with its contractions on the device:
    m <- m V and J <- J V (ONE contraction over all n * d_Q Jacobian rows), q <- (q - b) Phi, J <- Phi^T J (one MN-major
    contraction per sample).  Same dictionary contract as the host function: 'm_full' / 'q_full' / 'J_full' keep the
    unreduced arrays; the reduced 'm_data' / 'q_data' / 'J_data' are float32."""
    import torch
    assert input_pruning or output_pruning
    dicts = (train_dict, test_dict)
    key = 'J_data' if 'J_data' in train_dict else 'PhiTJ_data'
    already = key == 'PhiTJ_data'
    if key not in train_dict:
        raise KeyError('J_data')
    back = (lambda t: t.cpu().numpy()) if as_numpy else (lambda t: t)
    for d in dicts:
        d['J_full'] = d[key]
        Jd = _dev(d[key])
        if input_pruning:
            V = _dev(projector_dict['input'])                # [d_M, r_in]
            assert V.shape[0] == np.shape(d['m_data'])[1]
            Vt = V.t().contiguous()                          # [r_in, d_M]: K-major operand
            d['m_full'] = d['m_data']
            d['m_data'] = back(gemm_nt(_dev(d['m_data']), Vt).contiguous())
            n, dq, dM = Jd.shape
            Jd = gemm_nt(Jd.reshape(n * dq, dM), Vt).contiguous().reshape(n, dq, V.shape[1])
        if output_pruning:
            Phi, bias = _dev(projector_dict['output']), _dev(projector_dict['last_layer_bias'])   # [d_Q, r_out], [d_Q]
            assert Phi.shape[0] == np.shape(d['q_data'])[1]
            d['q_full'] = d['q_data']
            d['q_data'] = back(gemm_nt((_dev(d['q_data']) - bias).contiguous(), Phi.t().contiguous()).contiguous())
            if already:
                assert Jd.shape[1] == Phi.shape[1]
            else:
                Jd = torch.stack([gemm_tn(Phi, Jd[i].contiguous()) for i in range(Jd.shape[0])])
        d['J_data'] = back(Jd)
    return train_dict, test_dict


def upload_factors(data_dict):
    """m, q and the singular factors U, sigma, V of a dataset as device tensors (float32): uploaded ONCE, so that the
    per-epoch re-flattening of the sketched training (jacobianTrainingDriver.py:234-241) never goes back to the host."""
    return {k: _dev(data_dict[k]) for k in ('m_data', 'q_data', 'U_data', 'sigma_data', 'V_data')}


def flatten_data_device(data_dict, target_rank=80, batch_rank=8, order_random=True, diagonalize_sigma=True, seed=0,
                        burn_in=0):
    """flatten_data (dataUtilities.py:348-445) on device tensors: truncate the factors to target_rank, draw the same column
    permutation as the host function (numpy RandomState(seed) after burn_in draws), cut the columns into
    target_rank / batch_rank groups and make every group a training row [m, U_g, V_g] -> [q, diag(sigma_g)].  Pure data
    movement: the result equals the host function's cast to float32, element for element.  Inputs may be numpy arrays or
    the tensors of upload_factors; the returned dict holds device tensors, which DinoModel.fit / evaluate take as they
    are."""
    import torch
    assert target_rank % batch_rank == 0
    groups = target_rank // batch_rank
    m, q = _dev(data_dict['m_data']), _dev(data_dict['q_data'])
    U = _dev(data_dict['U_data'])[:, :, :target_rank]
    sg = _dev(data_dict['sigma_data'])[:, :target_rank]
    V = _dev(data_dict['V_data'])[:, :, :target_rank]
    if order_random:
        random_state = np.random.RandomState(seed=seed)
        for _ in range(burn_in):
            random_state.permutation(target_rank)
        cols = torch.from_numpy(random_state.permutation(target_rank).astype(np.int64)).to(U.device)
        U, sg, V = U.index_select(2, cols), sg.index_select(1, cols), V.index_select(2, cols)
    n, dQ, dM = m.shape[0], U.shape[1], V.shape[1]
    rows = n * groups
    Unew = U.reshape(n, dQ, groups, batch_rank).permute(0, 2, 1, 3).contiguous().reshape(rows, dQ, batch_rank)
    Vnew = V.reshape(n, dM, groups, batch_rank).permute(0, 2, 1, 3).contiguous().reshape(rows, dM, batch_rank)
    snew = sg.contiguous().reshape(rows, batch_rank)
    if diagonalize_sigma:
        snew = torch.diag_embed(snew)
    return {'m_data': m.repeat_interleave(groups, dim=0), 'q_data': q.repeat_interleave(groups, dim=0),
            'U_data': Unew, 'sigma_data': snew, 'V_data': Vnew}
