"""Data loading and target construction with the reference's names, on-disk formats and array
layouts (dataUtilities.py:21-548).  Host-side numpy: these run once per training run, before the
hot loop (SURVEY.md 8f rank 2 lists them as the next candidates for device kernels).
"""
import os
import re

import numpy as np


def _per_process_files(data_dir, pattern):
    """{process id: path} of the files in data_dir whose name matches ``pattern`` (one int group)."""
    found = {}
    for name in os.listdir(data_dir):
        hit = re.fullmatch(pattern, name)
        if hit:
            found[int(hit.group(1))] = data_dir + name
    return found


def _concat_in_process_order(files, loader):
    if not files:
        return None
    # the reference walks ranks 0..max and fails on a gap; keep that strictness
    parts = [loader(files[r]) for r in range(max(files) + 1)]
    return [np.concatenate([p[i] for p in parts]) for i in range(len(parts[0]))]


def load_data(data_dir, rescale=False, derivatives=False, n_data=None):
    """Concatenate the per-process sample files of ``data_dir`` (dataUtilities.py:21-153).

    (m, q): ``mq_on_proc{r}.npz`` with keys m_data/q_data, or legacy ``ms_on_proc_{r}.npy`` +
    ``qs_on_proc_{r}.npy``.  Jacobian factors: ``J_on_proc{r}.npz`` with keys
    U_data/sigma_data/V_data, or legacy ``Us_/sigmas_/Vs_on_proc_{r}.npy``."""
    assert os.path.isdir(data_dir)
    npz = _per_process_files(data_dir, r'mq_on_proc(\d+)\.npz')
    if npz:
        def read(path):
            z = np.load(path)
            return z['m_data'], z['q_data']
        m_data, q_data = _concat_in_process_order(npz, read)
    else:
        ms = _per_process_files(data_dir, r'ms_on_proc_(\d+)\.npy')
        (m_data,) = _concat_in_process_order(ms, lambda p: (np.load(p),))
        (q_data,) = _concat_in_process_order(ms, lambda p: (np.load(p.replace('ms_on_proc_', 'qs_on_proc_')),))
    if n_data is not None:
        assert type(n_data) is int
        assert n_data <= m_data.shape[0], 'Requesting too much data, available number: ' + str(m_data.shape[0]) + ', try again'
        m_data, q_data = m_data[:n_data], q_data[:n_data]
    if rescale:
        from sklearn import preprocessing
        m_data, q_data = preprocessing.scale(m_data), preprocessing.scale(q_data)
    data_dict = {'m_data': m_data, 'q_data': q_data}
    if derivatives:
        jnpz = _per_process_files(data_dir, r'J_on_proc(\d+)\.npz')
        if jnpz:
            def readj(path):
                z = np.load(path)
                return z['U_data'], z['sigma_data'], z['V_data']
            U_data, sigma_data, V_data = _concat_in_process_order(jnpz, readj)
        else:
            sig = _per_process_files(data_dir, r'sigmas_on_proc_(\d+)\.npy')
            def readl(path):
                return (np.load(path.replace('sigmas_on_proc_', 'Us_on_proc_')), np.load(path),
                        np.load(path.replace('sigmas_on_proc_', 'Vs_on_proc_')))
            U_data, sigma_data, V_data = _concat_in_process_order(sig, readl)
        if n_data is not None:
            U_data, sigma_data, V_data = U_data[:n_data], sigma_data[:n_data], V_data[:n_data]
        if rescale:
            raise NotImplementedError('This needs to be thought out with care')
        data_dict.update(U_data=U_data, sigma_data=sigma_data, V_data=V_data)
    return data_dict


def _load_basis(data_dir, files, exact, fragment, fixed_rank, spectrum_file, tolerance):
    name = exact if exact in files else next((f for f in files if 'AS' in f and fragment in f), None)
    basis = np.load(data_dir + name)
    if fixed_rank > 0:
        return basis[:, :fixed_rank]
    spectrum = np.load(data_dir + spectrum_file)
    return basis[:, np.where(spectrum > tolerance)[0]]


def get_projectors(data_dir, as_input_tolerance=1e-4, as_output_tolerance=1e-4, kle_tolerance=1e-4,
                   pod_tolerance=1e-4, fixed_input_rank=0, fixed_output_rank=0, mixed_output=True, verbose=False):
    """Reduced-basis files of a data directory (dataUtilities.py:156-244): AS_input_projector.npy /
    AS_output_projector.npy (mandatory; truncated to a fixed rank or by AS_d_GN.npy / AS_d_NG.npy),
    KLE_projector.npy and POD_projector.npy (optional).  Key names as in the reference, including
    its 'KLE' (not 'KLE_input') key."""
    files = os.listdir(data_dir)
    projectors = {}
    projectors['AS_input'] = _load_basis(data_dir, files, 'AS_input_projector.npy', 'input_projector.npy',
                                         fixed_input_rank, 'AS_d_GN.npy', as_input_tolerance)
    projectors['AS_output'] = _load_basis(data_dir, files, 'AS_output_projector.npy', 'output_projector.npy',
                                          fixed_output_rank, 'AS_d_NG.npy', as_output_tolerance)
    for key, fname, dname, rank, tol, msg in (
            ('KLE', 'KLE_projector.npy', 'KLE_d.npy', fixed_input_rank, kle_tolerance, 'KLE did not load'),
            ('POD', 'POD_projector.npy', 'POD_d.npy', fixed_output_rank, pod_tolerance, 'Pre-computed POD did not load')):
        try:
            basis = np.load(data_dir + fname)
            projectors[key] = basis[:, :rank] if rank > 0 else basis[:, np.where(np.load(data_dir + dname) > tol)[0]]
        except Exception:
            print(msg)
    if verbose:
        for k, v in projectors.items():
            print(k, 'projector shape = ', v.shape)
    return projectors


def _split_indices(random_indices, train_size, val_size, test_size, test_from_end):
    test = random_indices[-test_size:] if test_from_end else random_indices[:test_size]
    rest = random_indices[test_size:][:train_size + val_size]
    return rest[:train_size], rest[train_size:], test


def _shuffle_dict(m, q, tr, va, te, copy=False):
    c = np.copy if copy else (lambda a: a)
    return {'input_train': c(m[tr]), 'output_train': c(q[tr]), 'input_val': c(m[va]), 'output_val': c(q[va]),
            'input_test': c(m[te]), 'output_test': c(q[te])}


def shuffle_single_data(data_dictionary, train_size, val_size, test_size, seed=0, copy=True, burn_in=0):
    """One random train/val/test split of (m, q) (dataUtilities.py:252-292; the RNG seed is fixed to 0
    there regardless of ``seed``)."""
    assert type(train_size) is int and type(val_size) is int and type(test_size) is int
    m, q = data_dictionary['m_data'], data_dictionary['q_data']
    total = m.shape[0]
    assert train_size + val_size + test_size <= total
    random_state = np.random.RandomState(seed=0)
    for _ in range(burn_in):
        random_state.permutation(total)
    tr, va, te = _split_indices(random_state.permutation(total), train_size, val_size, test_size, False)
    return _shuffle_dict(m, q, tr, va, te)


def shuffle_data(data_dictionary, train_size, val_size, test_size, n_shuffles=10, seed=0, copy=True, burn_in=0):
    """n_shuffles random splits (dataUtilities.py:295-345); test indices come from the END of each draw."""
    assert type(train_size) is int and type(val_size) is int and type(test_size) is int
    m, q = data_dictionary['m_data'], data_dictionary['q_data']
    total = m.shape[0]
    assert train_size + val_size + test_size <= total
    if train_size + val_size + test_size == total:
        assert n_shuffles == 1
    random_state = np.random.RandomState(seed=0)
    for _ in range(burn_in):
        random_state.permutation(total)
    out = {}
    for shuffle in range(n_shuffles):
        tr, va, te = _split_indices(random_state.permutation(total), train_size, val_size, test_size, True)
        out[shuffle] = _shuffle_dict(m, q, tr, va, te, copy)
    return out


def flatten_data(data_dict, target_rank=80, batch_rank=8, order_random=True, diagonalize_sigma=True,
                 key_map={'m_data': 'm_data', 'q_data': 'q_data', 'U_data': 'U_data', 'sigma_data': 'sigma_data',
                          'V_data': 'V_data'},
                 seed=0, burn_in=0, verbose=False, independent_sampling=False):
    """Sketch minibatches for the matrix-subsampled truncated-SVD loss (dataUtilities.py:348-445).

    The rank-``target_rank`` factors of every sample are (optionally column-permuted and) cut into
    ``target_rank/batch_rank`` groups of ``batch_rank`` columns; each group becomes one training
    row [m, U_g, V_g] -> [q, diag(sigma_g)], so the output has n*target_rank/batch_rank rows."""
    assert target_rank % batch_rank == 0
    if independent_sampling:
        # the reference branch reads sigma_data_ before assigning it (dataUtilities.py:364, NameError)
        raise NotImplementedError('independent_sampling is broken in the reference and is not provided')
    groups = target_rank // batch_rank
    m_data, q_data = data_dict[key_map['m_data']], data_dict[key_map['q_data']]
    U = data_dict[key_map['U_data']][:, :, :target_rank]
    sigma = data_dict[key_map['sigma_data']][:, :target_rank]
    V = data_dict[key_map['V_data']][:, :, :target_rank]
    if order_random:
        random_state = np.random.RandomState(seed=seed)
        for _ in range(burn_in):
            random_state.permutation(target_rank)
        cols = random_state.permutation(target_rank)
        U, sigma, V = U[:, :, cols], sigma[:, cols], V[:, :, cols]
    n, dQ, dM = m_data.shape[0], U.shape[1], V.shape[1]
    rows = n * groups
    if verbose:
        print('n_data = ', n, ' n_batch_data = ', rows, ' dM = ', dM, ' dQ = ', dQ)
    # [n, d, groups*k] -> [n*groups, d, k]
    Unew = np.ascontiguousarray(U.reshape(n, dQ, groups, batch_rank).transpose(0, 2, 1, 3)).reshape(rows, dQ, batch_rank)
    Vnew = np.ascontiguousarray(V.reshape(n, dM, groups, batch_rank).transpose(0, 2, 1, 3)).reshape(rows, dM, batch_rank)
    sigmanew = sigma.reshape(rows, batch_rank)
    if diagonalize_sigma:
        diag = np.zeros((rows, batch_rank, batch_rank))
        d = np.arange(batch_rank)
        diag[:, d, d] = sigmanew
        sigmanew = diag
    return {key_map['m_data']: np.repeat(m_data, groups, axis=0), key_map['q_data']: np.repeat(q_data, groups, axis=0),
            key_map['U_data']: Unew, key_map['sigma_data']: sigmanew, key_map['V_data']: Vnew}


def train_test_split(data_dict, n_train, seed=0):
    """Deterministic split: the LAST n_train rows train, the rest test (dataUtilities.py:448-489)."""
    keys = ['m_data', 'q_data'] + (['U_data', 'sigma_data', 'V_data'] if 'U_data' in data_dict else [])
    train_dict = {k: data_dict[k][-n_train:] for k in keys}
    test_dict = {k: data_dict[k][:-n_train] for k in keys}
    return train_dict, test_dict


def _dense_jacobians(U_data, sigma_data, V_data):
    # J[i] = U[i] diag(sigma[i]) V[i]^T, fp64 [n, d_Q, d_M]
    return np.matmul(U_data * sigma_data[:, None, :], np.transpose(V_data, (0, 2, 1)))


def remap_jacobian_data(data_dict):
    """Dense Jacobian targets from their singular factors (dataUtilities.py:492-510)."""
    return {'m_data': data_dict['m_data'], 'q_data': data_dict['q_data'],
            'J_data': _dense_jacobians(data_dict['U_data'], data_dict['sigma_data'], data_dict['V_data'])}


def remap_jacobian_data_with_orth_V(data_dict):
    """As remap_jacobian_data, plus the QR-orthonormalised right factors and J V V^T built from the
    ORIGINAL V (dataUtilities.py:512-548)."""
    V_data = data_dict['V_data']
    new_dict = remap_jacobian_data(data_dict)
    new_dict['V_data'] = np.stack([np.linalg.qr(V_data[i])[0] for i in range(V_data.shape[0])])
    JV = np.matmul(new_dict['J_data'], V_data)
    new_dict['JVVT_data'] = np.matmul(JV, np.transpose(V_data, (0, 2, 1)))
    return new_dict
