"""CPU oracle for the DINO derivative-informed (H1) training step.

TEST INFRASTRUCTURE ONLY.  Nothing under ``dino_b200/`` may import this file.  It is used by
``tests/``, by ``__graft_entry__.smoke()`` as the checker and by ``bench.py`` for the
``cpu_baseline`` / ``--impl reference`` legs.  It is never the thing shipped or measured as
the product.

PARITY STATUS: **partially pinned**.  The reference (tomoleary/dino) holds no tests, golden
vectors or fixtures, and its arithmetic lives in TensorFlow == 2.7.0 (un-vendored, pinned only
in the reference's INSTALL.md:22,33; not installable here: no cp312 wheel, no network).

  * pinned against the reference's own code run in this container (tests/golden/make_golden.py
    executes the reference functions, TF elementwise ops replaced by numpy stand-ins):
    the four loss/metric formulas (trainingUtilities.py:25-51), ``prune_the_data``
    (jacobianTrainingDriver.py:93-185), ``remap_jacobian_data`` / ``flatten_data`` /
    ``train_test_split`` (dataUtilities.py:348-510).
  * UNPINNED (restated from TensorFlow/Keras 2.7 published semantics): Dense = xW+b with
    kernel layout [in,out]; softplus / SoftplusGrad; GradientTape.batch_jacobian; tf.gradients;
    Keras ``fit``/``evaluate`` batching; Keras Adam (epsilon outside the bias correction).

The Jacobian is computed the reference's way - d_out reverse passes through autograd with
``create_graph=True`` (jacobianConstruction.py:145-149 uses tape.batch_jacobian, which TF
vectorises into d_out vector-Jacobian products) - and the weight gradient is autograd's
double-backward, so none of the hand-derived chain formulas the CUDA kernels implement are
used here.  That independence is the point.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

ACT_LINEAR = 0
ACT_SOFTPLUS = 1


# --------------------------------------------------------------------------------------
# network description (follows neuralNetworks.py:21-73, 77-89, 93-172)
# --------------------------------------------------------------------------------------
@dataclass
class OLayer:
    name: str
    kernel: torch.Tensor                 # [in, out]   (Keras layout, y = x W + b)
    bias: Optional[torch.Tensor]         # [out] or None (use_bias=False)
    act: int = ACT_LINEAR
    trainable: bool = True
    # index of the layer whose INPUT is added to this layer's output (ResNet skip), or -1
    skip_from: int = -1


@dataclass
class ONet:
    d_in: int
    layers: List[OLayer] = field(default_factory=list)

    @property
    def d_out(self) -> int:
        return self.layers[-1].kernel.shape[1]

    def trainable_tensors(self) -> List[torch.Tensor]:
        out = []
        for L in self.layers:
            if L.trainable:
                out.append(L.kernel)
                if L.bias is not None:
                    out.append(L.bias)
        return out

    def to(self, dtype) -> "ONet":
        new = ONet(self.d_in)
        for L in self.layers:
            new.layers.append(OLayer(L.name, L.kernel.detach().to(dtype).clone(),
                                     None if L.bias is None else L.bias.detach().to(dtype).clone(),
                                     L.act, L.trainable, L.skip_from))
        return new

    def get_weights(self) -> Dict[str, List[np.ndarray]]:
        d = {}
        for L in self.layers:
            w = [L.kernel.detach().numpy().copy()]
            if L.bias is not None:
                w.append(L.bias.detach().numpy().copy())
            d[L.name] = w
        return d


def glorot_uniform(rng: np.random.Generator, fan_in: int, fan_out: int) -> np.ndarray:
    """Keras default Dense kernel initialiser (TF-knowledge): U(-l, l), l = sqrt(6/(in+out))."""
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=(fan_in, fan_out))


def _t(a, dtype):
    return torch.tensor(np.asarray(a), dtype=dtype)


def build_projected_dense(rng, *, input_projector=None, last_layer_weights=None,
                          hidden_layer_dimensions: Sequence[int], compat_layer=True,
                          reduced_input_dim=None, reduced_output_dim=None,
                          name_prefix="", dtype=torch.float64,
                          first_layer_trainable=False, last_layer_trainable=False) -> ONet:
    """Restates projected_dense (neuralNetworks.py:21-73)."""
    layers: List[OLayer] = []
    if input_projector is None:
        d_in = reduced_input_dim
    else:
        d_in, reduced_input_dim = input_projector.shape
        layers.append(OLayer(name_prefix + "input_proj_layer", _t(input_projector, dtype), None,
                             ACT_LINEAR, first_layer_trainable))
    if last_layer_weights is not None:
        reduced_output_dim, output_dim = last_layer_weights[0].shape
    assert hidden_layer_dimensions[-1] == reduced_output_dim
    k = reduced_input_dim
    layers.append(OLayer(name_prefix + "dense_reduction_layer",
                         _t(glorot_uniform(rng, k, reduced_input_dim), dtype),
                         torch.zeros(reduced_input_dim, dtype=dtype), ACT_SOFTPLUS))
    for i, h in enumerate(hidden_layer_dimensions):
        layers.append(OLayer(name_prefix + "inner_layer_" + str(i), _t(glorot_uniform(rng, k, h), dtype),
                             torch.zeros(h, dtype=dtype), ACT_SOFTPLUS))
        k = h
    if compat_layer:
        layers.append(OLayer(name_prefix + "output_compat_layer",
                             _t(glorot_uniform(rng, k, reduced_output_dim), dtype), None, ACT_LINEAR))
        k = reduced_output_dim
    if last_layer_weights is not None:
        layers.append(OLayer(name_prefix + "output_layer", _t(last_layer_weights[0], dtype),
                             _t(last_layer_weights[1], dtype), ACT_LINEAR, last_layer_trainable))
    return ONet(d_in, layers)


def build_generic_dense(rng, input_dim, output_dim, n_hidden_neurons, name_prefix="",
                        dtype=torch.float64) -> ONet:
    """Restates generic_dense (neuralNetworks.py:77-89)."""
    layers, k = [], input_dim
    for i, h in enumerate(n_hidden_neurons):
        layers.append(OLayer(name_prefix + "inner_layer_" + str(i), _t(glorot_uniform(rng, k, h), dtype),
                             torch.zeros(h, dtype=dtype), ACT_SOFTPLUS))
        k = h
    layers.append(OLayer(name_prefix + "final_dense", _t(glorot_uniform(rng, k, output_dim), dtype),
                         torch.zeros(output_dim, dtype=dtype), ACT_LINEAR))
    return ONet(input_dim, layers)


def build_projected_resnet(rng, *, input_projector=None, last_layer_weights=None, ranks: Sequence[int],
                           compat_layer=True, reduced_input_dim=None, reduced_output_dim=None,
                           name_prefix="", dtype=torch.float64, first_layer_trainable=False,
                           last_layer_trainable=True, zeros=True) -> ONet:
    """The INTENDED projected_resnet (neuralNetworks.py:122-172; as shipped it raises NameError
    at :142).  Block i: z <- z + softplus(z W1 + b1) W2 + b2, W2/b2 zero-initialised (:112-114)."""
    layers: List[OLayer] = []
    if input_projector is None:
        d_in = reduced_input_dim
    else:
        d_in, reduced_input_dim = input_projector.shape
        layers.append(OLayer(name_prefix + "input_proj_layer", _t(input_projector, dtype), None,
                             ACT_LINEAR, first_layer_trainable))
    if last_layer_weights is not None:
        reduced_output_dim, output_dim = last_layer_weights[0].shape
    k = reduced_input_dim
    if reduced_input_dim != reduced_output_dim:
        layers.append(OLayer(name_prefix + "input_resnet_interface",
                             _t(glorot_uniform(rng, k, reduced_output_dim), dtype), None, ACT_LINEAR))
        k = reduced_output_dim
    for i, r in enumerate(ranks):
        layers.append(OLayer(name_prefix + str(i) + "low_rank_residual_in", _t(glorot_uniform(rng, k, r), dtype),
                             torch.zeros(r, dtype=dtype), ACT_SOFTPLUS))
        w2 = np.zeros((r, k)) if zeros else glorot_uniform(rng, r, k)
        layers.append(OLayer(name_prefix + str(i) + "low_rank_residual_out", _t(w2, dtype),
                             torch.zeros(k, dtype=dtype), ACT_LINEAR, True, skip_from=len(layers) - 1))
    if compat_layer:
        layers.append(OLayer(name_prefix + "output_compat_layer",
                             _t(glorot_uniform(rng, k, reduced_output_dim), dtype), None, ACT_LINEAR))
    if last_layer_weights is not None:
        layers.append(OLayer(name_prefix + "output_layer", _t(last_layer_weights[0], dtype),
                             _t(last_layer_weights[1], dtype), ACT_LINEAR, last_layer_trainable))
    return ONet(d_in, layers)


# --------------------------------------------------------------------------------------
# forward, Jacobians (jacobianConstruction.py:102-185)
# --------------------------------------------------------------------------------------
def softplus(x: torch.Tensor) -> torch.Tensor:
    # log(1+e^x), stable.  TF's kernel switches to x / e^x outside +-13.94 (fp32): the difference
    # is below fp32 epsilon of the result.
    return torch.nn.functional.softplus(x, beta=1.0, threshold=1e9)


def forward(net: ONet, m: torch.Tensor) -> torch.Tensor:
    inputs = []
    z = m
    for L in net.layers:
        inputs.append(z)
        a = z @ L.kernel
        if L.bias is not None:
            a = a + L.bias
        if L.act == ACT_SOFTPLUS:
            a = softplus(a)
        if L.skip_from >= 0:
            a = a + inputs[L.skip_from]
        z = a
    return z


def batch_jacobian(net: ONet, m: torch.Tensor, create_graph: bool, vectorised: bool = False):
    """q_hat [B,d_out] and J_hat[b,i,j] = d q_hat[b,i] / d m[b,j]  ([B,d_out,d_in]) obtained the
    reference's way: one reverse pass per output component (equip_model_with_full_jacobian,
    jacobianConstruction.py:128-153)."""
    m = m.detach().requires_grad_(True)
    q = forward(net, m)
    d_out = q.shape[1]
    if vectorised:
        eye = torch.eye(d_out, dtype=q.dtype).unsqueeze(1).expand(d_out, q.shape[0], d_out)
        (g,) = torch.autograd.grad(q, m, grad_outputs=eye, is_grads_batched=True,
                                   create_graph=create_graph, retain_graph=True)
        J = g.transpose(0, 1)
    else:
        rows = []
        for i in range(d_out):
            (g,) = torch.autograd.grad(q[:, i].sum(), m, create_graph=create_graph, retain_graph=True)
            rows.append(g)
        J = torch.stack(rows, dim=1)
    return q, J


def output_reduced_jacobian(net: ONet, m, Phi: torch.Tensor, create_graph: bool):
    """q_hat and Phi^T J_hat  [B, r_out, d_in]  (equip_model_with_output_reduced_jacobian,
    jacobianConstruction.py:155-185: einsum('ij,kil->kjl', Phi, J))."""
    q, J = batch_jacobian(net, m, create_graph)
    return q, torch.einsum("ij,kil->kjl", Phi, J)


def sketched_jacobian(net: ONet, m, U: torch.Tensor, V: torch.Tensor, create_graph: bool):
    """q_hat and U^T J_hat V  [B,k,k]  (equip_model_with_sketched_jacobian,
    jacobianConstruction.py:102-125): k reverse passes seeded with the columns of U."""
    m = m.detach().requires_grad_(True)
    q = forward(net, m)
    UTy = torch.einsum("ijk,ij->ik", U, q)
    rows = []
    for i in range(UTy.shape[1]):
        (g,) = torch.autograd.grad(UTy[:, i].sum(), m, create_graph=create_graph, retain_graph=True)
        rows.append(g)
    UTdydx = torch.stack(rows, dim=1)                       # [B,k,d_M]
    return q, torch.einsum("ijk,ikl->ijl", UTdydx, V)


# --------------------------------------------------------------------------------------
# losses and metrics (trainingUtilities.py:25-51)
# --------------------------------------------------------------------------------------
def normalized_mse(y_true, y_pred):
    sq = (y_true - y_pred) ** 2
    return (sq.mean(-1) / (y_true ** 2).mean(-1)).mean()


def normalized_mse_matrix(y_true, y_pred):
    sq = (y_true - y_pred) ** 2
    return (sq.sum(-1).sum(-1) / (y_true ** 2).sum(-1).sum(-1)).mean()


def l2_accuracy(y_true, y_pred):
    sq = (y_true - y_pred) ** 2
    return 1.0 - torch.sqrt((sq.mean(-1) / (y_true ** 2).mean(-1)).mean())


def f_accuracy_matrix(y_true, y_pred):
    sq = (y_true - y_pred) ** 2
    return 1.0 - torch.sqrt((sq.sum(-1).sum(-1) / (y_true ** 2).sum(-1).sum(-1)).mean())


@dataclass
class Batch:
    m: torch.Tensor
    q: torch.Tensor
    J: torch.Tensor                      # full / out_reduced: [B,s,d_in]; sketched: [B,k,k] (diag sigma)
    U: Optional[torch.Tensor] = None     # sketched only [B,d_Q,k]
    V: Optional[torch.Tensor] = None     # sketched only [B,d_M,k]


def model_outputs(net: ONet, batch: Batch, jac_mode: str, create_graph: bool, Phi=None, vectorised=False):
    if jac_mode == "full":
        return batch_jacobian(net, batch.m, create_graph, vectorised)
    if jac_mode == "out_reduced":
        return output_reduced_jacobian(net, batch.m, Phi, create_graph)
    if jac_mode == "sketched":
        return sketched_jacobian(net, batch.m, batch.U, batch.V, create_graph)
    raise ValueError(jac_mode)


def h1_loss(net: ONet, batch: Batch, loss_weights=(1.0, 1.0), jac_mode="full", Phi=None,
            create_graph=True, vectorised=False):
    """Total Keras loss w0*normalized_mse(q) + w1*normalized_mse_matrix(J) (trainingUtilities.py:99-102)."""
    qh, Jh = model_outputs(net, batch, jac_mode, create_graph, Phi, vectorised)
    l2 = normalized_mse(batch.q, qh)
    h1 = normalized_mse_matrix(batch.J, Jh)
    return loss_weights[0] * l2 + loss_weights[1] * h1, l2, h1, qh, Jh


def evaluate_batch(net: ONet, batch: Batch, loss_weights=(1.0, 1.0), jac_mode="full", Phi=None):
    """The 7 numbers Keras ``evaluate`` reports for ONE batch: [loss, l2, h1, l2_accuracy(q),
    f_accuracy_matrix(q), l2_accuracy(J), f_accuracy_matrix(J)] (metrics list applied to both
    outputs; index map: surrogatePostProcessing.py:138)."""
    with torch.enable_grad():
        total, l2, h1, qh, Jh = h1_loss(net, batch, loss_weights, jac_mode, Phi, create_graph=False)
    qh, Jh = qh.detach(), Jh.detach()
    vals = [total.detach(), l2.detach(), h1.detach(), l2_accuracy(batch.q, qh), f_accuracy_matrix(batch.q, qh),
            l2_accuracy(batch.J, Jh), f_accuracy_matrix(batch.J, Jh)]
    return [float(v) for v in vals]


def slice_batch(data: Batch, idx) -> Batch:
    idx = torch.as_tensor(np.asarray(idx), dtype=torch.long)
    return Batch(data.m[idx], data.q[idx], data.J[idx],
                 None if data.U is None else data.U[idx], None if data.V is None else data.V[idx])


def evaluate(net: ONet, data: Batch, batch_size=32, **kw):
    """Keras ``evaluate`` (TF-knowledge, Keras 2.7): the three losses are averaged with batch-size weights
    (LossesContainer: sample_weight = batch size); the four function metrics return one scalar per batch and go
    through MeanMetricWrapper, so each batch counts once, whatever its size."""
    n = data.m.shape[0]
    loss_acc, met_acc, nb = np.zeros(3), np.zeros(4), 0
    for s in range(0, n, batch_size):
        idx = np.arange(s, min(n, s + batch_size))
        vals = np.array(evaluate_batch(net, slice_batch(data, idx), **kw))
        loss_acc += vals[:3] * len(idx)
        met_acc += vals[3:]
        nb += 1
    return list(loss_acc / n) + list(met_acc / nb)


# --------------------------------------------------------------------------------------
# optimisers (Keras semantics, TF-knowledge; selected at trainingUtilities.py:89-94)
# --------------------------------------------------------------------------------------
class KerasAdam:
    """lr_t = lr*sqrt(1-b2^t)/(1-b1^t); m = b1 m+(1-b1) g; v = b2 v+(1-b2) g^2;
    theta -= lr_t*m/(sqrt(v)+eps), eps = 1e-7 OUTSIDE the bias correction."""

    def __init__(self, params, lr=1e-3, b1=0.9, b2=0.999, eps=1e-7):
        self.params = list(params)
        self.lr, self.b1, self.b2, self.eps = lr, b1, b2, eps
        self.m = [torch.zeros_like(p) for p in self.params]
        self.v = [torch.zeros_like(p) for p in self.params]
        self.t = 0

    def step(self, grads):
        self.t += 1
        lr_t = self.lr * math.sqrt(1.0 - self.b2 ** self.t) / (1.0 - self.b1 ** self.t)
        with torch.no_grad():
            for p, g, m, v in zip(self.params, grads, self.m, self.v):
                m.mul_(self.b1).add_(g, alpha=1.0 - self.b1)
                v.mul_(self.b2).addcmul_(g, g, value=1.0 - self.b2)
                p.sub_(lr_t * m / (v.sqrt() + self.eps))


class KerasSGD:
    def __init__(self, params, lr=1e-2):
        self.params, self.lr = list(params), lr

    def step(self, grads):
        with torch.no_grad():
            for p, g in zip(self.params, grads):
                p.sub_(self.lr * g)


def h1_step_grads(net: ONet, batch: Batch, loss_weights=(1.0, 1.0), jac_mode="full", Phi=None,
                  vectorised=False):
    """Loss and d loss / d theta for the trainable tensors, by autograd double-backward."""
    params = net.trainable_tensors()
    for p in params:
        p.requires_grad_(True)
    total, l2, h1, _, _ = h1_loss(net, batch, loss_weights, jac_mode, Phi, True, vectorised)
    grads = torch.autograd.grad(total, params, allow_unused=True)
    grads = [torch.zeros_like(p) if g is None else g for p, g in zip(params, grads)]
    for p in params:
        p.requires_grad_(False)
    return (float(total.detach()), float(l2.detach()), float(h1.detach())), grads


def fit(net: ONet, data: Batch, *, epochs=1, batch_size=32, lr=1e-3, opt="adam",
        loss_weights=(1.0, 1.0), jac_mode="full", Phi=None, perms: Optional[List[np.ndarray]] = None,
        vectorised=False, optimizer=None, max_steps=None):
    """Keras ``fit`` semantics (trainingUtilities.py:135-138): each epoch walks the permutation
    in chunks of batch_size, the last chunk may be short and is averaged over its own size.
    ``perms[e]`` is the sample order of epoch e (identity if None: shuffle=False)."""
    n = data.m.shape[0]
    if optimizer is None:
        optimizer = KerasAdam(net.trainable_tensors(), lr) if opt == "adam" else KerasSGD(net.trainable_tensors(), lr)
    history, steps = [], 0
    for e in range(epochs):
        order = np.arange(n) if perms is None else np.asarray(perms[e])
        for s in range(0, n, batch_size):
            idx = order[s:s + batch_size]
            losses, grads = h1_step_grads(net, slice_batch(data, idx), loss_weights, jac_mode, Phi, vectorised)
            optimizer.step(grads)
            history.append(losses)
            steps += 1
            if max_steps is not None and steps >= max_steps:
                return history, optimizer
    return history, optimizer


# --------------------------------------------------------------------------------------
# data-side restatements (dataUtilities.py, jacobianTrainingDriver.py:93-185)
# --------------------------------------------------------------------------------------
def remap_jacobian(U, sigma, V):
    """J[i] = U[i] (sigma[i] * V[i])^T   (remap_jacobian_data, dataUtilities.py:492-510)."""
    return np.einsum("nqr,nr,nmr->nqm", U, sigma, V)


def prune(projector_in, projector_out, last_layer_bias, m, q, J, input_pruning=True, output_pruning=True):
    """Reduced-coordinate targets (prune_the_data, jacobianTrainingDriver.py:134,143,167,173)."""
    if input_pruning:
        m = m @ projector_in
        J = J @ projector_in
    if output_pruning:
        q = (q - last_layer_bias) @ projector_out
        J = np.einsum("ji,kjl->kil", projector_out, J)
    return m, q, J


def flatten(m, q, U, sigma, V, target_rank, batch_rank, order_random=True, seed=0, burn_in=0):
    """Sketch minibatches (flatten_data, dataUtilities.py:348-445, default branch)."""
    assert target_rank % batch_rank == 0
    bf = target_rank // batch_rank
    U, sigma, V = U[:, :, :target_rank], sigma[:, :target_rank], V[:, :, :target_rank]
    if order_random:
        rs = np.random.RandomState(seed=seed)
        for _ in range(burn_in):
            rs.permutation(target_rank)
        p = rs.permutation(target_rank)
        U, sigma, V = U[:, :, p], sigma[:, p], V[:, :, p]
    n, dQ, dM = m.shape[0], U.shape[1], V.shape[1]
    nb = n * bf
    mnew, qnew = np.repeat(m, bf, axis=0), np.repeat(q, bf, axis=0)
    Unew = U.transpose(0, 2, 1).reshape(nb, batch_rank, dQ).transpose(0, 2, 1)
    Vnew = V.transpose(0, 2, 1).reshape(nb, batch_rank, dM).transpose(0, 2, 1)
    sg = sigma.reshape(nb, batch_rank)
    S = np.zeros((nb, batch_rank, batch_rank))
    S[:, np.arange(batch_rank), np.arange(batch_rank)] = sg
    return mnew, qnew, Unew, S, Vnew


def jacobian_errors(J_pred: np.ndarray, J_true: np.ndarray) -> Dict[str, float]:
    """Per-sample relative errors of compute_jacobian_errors (jacobianTests.py:104-126)."""
    n = np.linalg.norm
    U, _, VT = np.linalg.svd(J_true, full_matrices=False)
    out = {
        "Jerror": n(J_pred - J_true) / n(J_true),
        "JTJerror": n(J_pred.T @ J_pred - J_true.T @ J_true) / n(J_true.T @ J_true),
        "JJTerror": n(J_pred @ J_pred.T - J_true @ J_true.T) / n(J_true @ J_true.T),
        "reducedJerror": n(U.T @ J_pred @ VT.T - U.T @ J_true @ VT.T) / n(U.T @ J_true @ VT.T),
        "nullspace_error": n(J_pred - J_pred @ VT.T @ VT) / n(J_true),
        "rJTJerror": n(VT @ J_pred.T @ J_pred @ VT.T - VT @ J_true.T @ J_true @ VT.T) / n(VT @ J_true.T @ J_true @ VT.T),
        "rJJTerror": n(U.T @ J_pred @ J_pred.T @ U - U.T @ J_true @ J_true.T @ U) / n(U.T @ J_true @ J_true.T @ U),
    }
    return {k: float(v) for k, v in out.items()}
