"""numpy model of the kernel pipeline (test infrastructure, NOT a product code path).

Restates, stage by stage, exactly what the CUDA step enqueues (dino_b200/csrc/plan.cu):
MLP forward -> reverse Jacobian chain -> losses -> chain backward -> MLP backward.  It exists so
the hand-derived double-backward formulas can be checked against the
autograd oracle on a machine without a GPU, and so a failing GPU test can be bisected by stage.
"""
import numpy as np

ACT_LINEAR, ACT_SOFTPLUS = 0, 1


def sigmoid(x):
    return 0.5 * (1.0 + np.tanh(0.5 * x))


def softplus(x):
    return np.maximum(x, 0.0) + np.log1p(np.exp(-np.abs(x)))


class Layer:
    def __init__(self, W, b, act, trainable=True, skip_from=-1):
        self.W, self.b, self.act, self.trainable, self.skip_from = W, b, act, trainable, skip_from


def mlp_forward(layers, m):
    z, s = [m], []
    for L in layers:
        a = z[-1] @ L.W
        if L.b is not None:
            a = a + L.b
        if L.act == ACT_SOFTPLUS:
            out, sl = softplus(a), sigmoid(a)
        else:
            out, sl = a, np.ones_like(a)
        if L.skip_from >= 0:
            out = out + z[L.skip_from]
        z.append(out)
        s.append(sl)
    return z, s


def reverse_chain(layers, s, seed):
    """seed [B,S,d_out] -> Q[l] = R[l+1]*s[l] for every layer and R0 = seeded Jacobian [B,S,d_in]."""
    nL = len(layers)
    Q = [None] * nL
    R_next = seed
    for l in range(nL - 1, -1, -1):
        Q[l] = R_next * s[l][:, None, :]
        R = Q[l] @ layers[l].W.T
        for l2 in range(l + 1, nL):                      # skips that land on the input of layer l
            if layers[l2].skip_from == l:
                assert layers[l2].act == ACT_LINEAR
                R = R + Q[l2]
        R_next = R
    return Q, R_next


def step(layers, m, q, Jt, w=(1.0, 1.0), seed=None, V=None, inv_B=None, implicit_first_layer=False):
    """Returns dict(loss3, grads[(dW, db)], qhat, Jhat).  seed: None -> identity (full Jacobian);
    [S,d_out] shared (Phi^T); or [B,S,d_out] per sample (U^T, then V [B,d_M,k] closes the sketch).
    implicit_first_layer: layer 0 of the backward chain and the H1 loss as csrc/plan.cu's run_implicit_layer0 computes
    them - from T = J W_1, M = W_1^T W_1 and the chain state Q above layer 0, never from J_hat (full Jacobian only)."""
    B = m.shape[0]
    inv_B = 1.0 / B if inv_B is None else inv_B
    nL = len(layers)
    d_out = layers[-1].W.shape[1]
    z, s = mlp_forward(layers, m)
    qh = z[-1]
    if seed is None:
        seed = np.eye(d_out)
    if seed.ndim == 2:
        seed = np.broadcast_to(seed, (B,) + seed.shape)
    Q, R0 = reverse_chain(layers, s, seed)
    Jh = R0 if V is None else R0 @ V
    # losses (trainingUtilities.py:25-37)
    eq, nq = ((q - qh) ** 2).sum(1), (q ** 2).sum(1)
    eJ, nJ = ((Jt - Jh) ** 2).sum((1, 2)), (Jt ** 2).sum((1, 2))
    if implicit_first_layer:
        assert V is None and layers[0].skip_from < 0
        W1, Q0 = layers[0].W, Q[0]
        cb = 2.0 * w[1] * inv_B / nJ                       # [B]
        T = Jt @ W1                                        # [B,S,n1]: the target is the operand of the d_in-wide contraction
        Mw = W1.T @ W1
        QM = Q0 @ Mw
        eJ = ((QM * Q0).sum(2) - 2.0 * (T * Q0).sum(2) + (Jt ** 2).sum(2)).sum(1)       # per-row identity, summed
    l2, h1 = (eq / nq).sum() * inv_B, (eJ / nJ).sum() * inv_B
    dq = 2.0 * w[0] * (qh - q) / nq[:, None] * inv_B
    dJ = 2.0 * w[1] * (Jh - Jt) / nJ[:, None, None] * inv_B
    dR = dJ if V is None else dJ @ V.transpose(0, 2, 1)
    # chain backward, ascending
    dW = [np.zeros_like(L.W) for L in layers]
    db = [None if L.b is None else np.zeros_like(L.b) for L in layers]
    rowsum = [None] * nL
    dR_in = {0: dR}
    for l in range(nL):
        if l == 0 and implicit_first_layer:
            Qc = cb[:, None, None] * Q0
            dQ = Qc @ Mw - cb[:, None, None] * T                                        # = G W_1
            if layers[0].trainable:
                dW[0] += W1 @ np.einsum("bsi,bso->io", Qc, Q0) - np.einsum("bsi,bso->io", Jt, Qc)   # = G^T Q
        else:
            dRl = dR_in[l]
            dQ = dRl @ layers[l].W
            if layers[l].trainable:
                dW[l] += np.einsum("bsi,bso->io", dRl, Q[l])
        rowsum[l] = (dQ * Q[l]).sum(1)
        nxt = dQ * s[l][:, None, :]
        if layers[l].skip_from >= 0:
            nxt = nxt + dR_in[layers[l].skip_from]
        dR_in[l + 1] = nxt
    # MLP backward, descending
    dz = {nL: dq}
    for l in range(nL - 1, -1, -1):
        L = layers[l]
        g = dz[l + 1]
        if L.act == ACT_SOFTPLUS:
            da = g * s[l] + rowsum[l] * (1.0 - s[l])
        else:
            da = g
        if L.trainable:
            dW[l] += z[l].T @ da
            if L.b is not None:
                db[l] += da.sum(0)
        dz[l] = dz.get(l, 0) + da @ L.W.T
        if L.skip_from >= 0:
            dz[L.skip_from] = dz.get(L.skip_from, 0) + g
    return dict(loss3=(w[0] * l2 + w[1] * h1, l2, h1), dW=dW, db=db, qhat=qh, Jhat=Jh)
