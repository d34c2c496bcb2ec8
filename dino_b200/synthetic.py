"""Synthetic (m, q, J) data and the BASELINE.json configurations.

The reference's data come from FEniCS/hIPPYlib PDE solves, which are out of scope; benchmarks and
tests use a random *teacher* network of the same architecture so that the regression problem is
learnable and every norm is O(1):  m ~ N(0, I),  q = teacher(m) + 0.01 N(0,1),
J = d teacher / d m + 0.01 N(0,1)/sqrt(d_in).  Generation is host-side numpy (it is data
preparation, not the training path).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np


def _softplus(x):
    return np.maximum(x, 0.0) + np.log1p(np.exp(-np.abs(x)))


def _sigmoid(x):
    return 0.5 * (1.0 + np.tanh(0.5 * x))


@dataclass
class Arch:
    """Architecture in reduced (training) coordinates."""
    kind: str                      # 'rb_dense' | 'rb_resnet' | 'generic_dense'
    d_in: int
    d_out: int
    hidden: List[int] = field(default_factory=list)   # rb_dense / generic_dense hidden widths
    ranks: List[int] = field(default_factory=list)    # rb_resnet block ranks
    frozen_output: bool = False    # trailing frozen (Phi^T, b) layer (reduced-input-only training)
    r_out: Optional[int] = None    # width in front of the output layer


CONFIGS: Dict[str, dict] = {
    # BASELINE.json configs[0]: rdiff, r_in = r_out = 50, reduced input and output
    "cfg1": dict(arch=Arch("rb_dense", 50, 50, hidden=6 * [50], r_out=50), n=1024, batch=32),
    # configs[1]: confusion, r_in = 100, d_Q = 50, reduced input only (frozen 50->50 output layer)
    "cfg2": dict(arch=Arch("rb_dense", 100, 50, hidden=6 * [50], frozen_output=True, r_out=50), n=4096, batch=256),
    # configs[2]: hyperelasticity, r_in = 200, r_out = 100, width 400
    "cfg3": dict(arch=Arch("rb_dense", 200, 100, hidden=5 * [400] + [100], frozen_output=True, r_out=100), n=8192,
                 batch=256),
    "cfg3b": dict(arch=Arch("rb_resnet", 200, 100, ranks=6 * [400], frozen_output=True, r_out=100), n=8192, batch=256),
    # configs[3]: full Jacobian, unreduced input
    "cfg4": dict(arch=Arch("generic_dense", 4225, 400, hidden=6 * [100]), n=512, batch=32),
    # configs[4]: scaling sweep, width 1024
    "cfg5": dict(arch=Arch("rb_dense", 200, 100, hidden=5 * [1024] + [100], frozen_output=True, r_out=100), n=65536,
                 batch=256),
}


def layer_table(arch: Arch):
    """[(name, n_in, n_out, softplus?, bias?, skip_from)] of the network in graph order."""
    rows, k = [], arch.d_in
    if arch.kind == "generic_dense":
        for i, h in enumerate(arch.hidden):
            rows.append((f"inner_layer_{i}", k, h, True, True, -1)); k = h
        rows.append(("final_dense", k, arch.d_out, False, True, -1))
        return rows
    r_out = arch.r_out or arch.d_out
    if arch.kind == "rb_dense":
        rows.append(("dense_reduction_layer", k, arch.d_in, True, True, -1)); k = arch.d_in
        for i, h in enumerate(arch.hidden):
            rows.append((f"inner_layer_{i}", k, h, True, True, -1)); k = h
    else:
        if arch.d_in != r_out:
            rows.append(("input_resnet_interface", k, r_out, False, False, -1)); k = r_out
        for i, r in enumerate(arch.ranks):
            rows.append((f"{i}low_rank_residual_in", k, r, True, True, -1))
            rows.append((f"{i}low_rank_residual_out", r, k, False, True, len(rows) - 1))
    rows.append(("output_compat_layer", k, r_out, False, False, -1)); k = r_out
    if arch.frozen_output:
        rows.append(("output_layer", k, arch.d_out, False, True, -1))
    return rows


def random_weights(arch: Arch, rng: np.random.Generator, teacher: bool):
    """{name: [kernel(, bias)]}: teacher = N(0, 1/fan_in) kernels, N(0, 0.1^2) biases; student =
    glorot-uniform kernels, zero biases (Keras defaults).  A frozen output layer is an orthonormal
    Phi^T with a small bias in both cases (the caller copies the teacher's into the student)."""
    out = {}
    for name, k, n, _sp, bias, _skip in layer_table(arch):
        if name == "output_layer":
            Phi, _ = np.linalg.qr(rng.normal(size=(n, k)))
            W = Phi.T if n >= k else np.linalg.qr(rng.normal(size=(k, n)))[0]
            b = rng.normal(0.0, 0.1, size=n)
        elif teacher:
            W = rng.normal(0.0, 1.0 / math.sqrt(k), size=(k, n))
            b = rng.normal(0.0, 0.1, size=n)
        else:
            lim = math.sqrt(6.0 / (k + n))
            W = rng.uniform(-lim, lim, size=(k, n))
            b = np.zeros(n)
        out[name] = [W.astype(np.float32)] + ([b.astype(np.float32)] if bias else [])
    return out


def student_weights(arch: Arch, seed: int, teacher_weights=None, bias_noise=0.0):
    """Student initialisation of SURVEY.md 8d: glorot-uniform kernels and zero biases from ``default_rng(seed)``; a
    frozen output layer is copied from the teacher; ``bias_noise`` > 0 draws N(0, bias_noise^2) biases instead (tests)."""
    rng = np.random.default_rng(seed)
    w = random_weights(arch, rng, teacher=False)
    if teacher_weights is not None and "output_layer" in w:
        w["output_layer"] = [a.copy() for a in teacher_weights["output_layer"]]
    if bias_noise:
        for name, arrs in w.items():
            if len(arrs) == 2 and name != "output_layer":
                arrs[1] = rng.normal(0, bias_noise, arrs[1].shape).astype(np.float32)
    return w


def teacher_eval(arch: Arch, weights, m: np.ndarray, want_jacobian=True, chunk=256):
    """q = net(m) and J = d net / d m ([n, d_out, d_in]) by a forward-mode chain, in float64."""
    table = layer_table(arch)
    n = m.shape[0]
    q = np.empty((n, arch.d_out))
    J = np.empty((n, arch.d_out, arch.d_in), dtype=np.float32) if want_jacobian else None
    for lo in range(0, n, chunk):
        z = m[lo:lo + chunk].astype(np.float64)
        T = None
        zs, Ts = [], []
        for name, k, h, sp, bias, skip in table:
            zs.append(z); Ts.append(T)
            W = weights[name][0].astype(np.float64)
            a = z @ W + (weights[name][1].astype(np.float64) if bias else 0.0)
            if want_jacobian:
                P = np.broadcast_to(W, (z.shape[0],) + W.shape) if T is None else T @ W
            if sp:
                z = _softplus(a)
                if want_jacobian:
                    P = P * _sigmoid(a)[:, None, :]
            else:
                z = a
            if skip >= 0:
                z = z + zs[skip]
                if want_jacobian:
                    P = P + (Ts[skip] if Ts[skip] is not None else np.eye(arch.d_in)[None])
            T = P if want_jacobian else None
        q[lo:lo + chunk] = z
        if want_jacobian:
            J[lo:lo + chunk] = np.transpose(T, (0, 2, 1))
    return q, J


def make_dataset(arch: Arch, n: int, seed: int, noise=0.01):
    """Teacher data of SURVEY.md 8d: returns dict(m, q, J) float32 plus the teacher weights."""
    rng = np.random.default_rng(seed)
    tw = random_weights(arch, rng, teacher=True)
    m = rng.normal(size=(n, arch.d_in))
    q, J = teacher_eval(arch, tw, m)
    q = q + noise * rng.normal(size=q.shape)
    J = J + (noise / math.sqrt(arch.d_in)) * rng.normal(size=J.shape).astype(np.float32)
    return dict(m=m.astype(np.float32), q=q.astype(np.float32), J=J.astype(np.float32)), tw


def make_device_dataset(arch: Arch, n: int, seed: int, noise=0.01, chunk=512):
    """The teacher data of ``make_dataset`` for large n / wide networks, produced ON the device: the teacher's outputs
    and Jacobians come from this library's own forward + Jacobian-chain kernels (fp32x3 mode), the noise from torch's
    device generator.  Data preparation, not the training path.  Returns (DeviceDataset, teacher weights, m on the host)."""
    import torch
    from .model import DeviceDataset
    from . import _lib
    rng = np.random.default_rng(seed)
    tw = random_weights(arch, rng, teacher=True)
    m_host = rng.standard_normal((n, arch.d_in), dtype=np.float32)
    teacher = build_model(arch, tw, precision="fp32x3")
    dev = teacher.core.device_params().device
    m = torch.from_numpy(m_host).to(dev)
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed + 7)
    q = torch.empty(n, arch.d_out, dtype=torch.float32, device=dev)
    J = torch.empty(n, arch.d_out, arch.d_in, dtype=torch.float32, device=dev)
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        part = DeviceDataset.from_device(_lib.JAC_FULL, m[lo:hi])
        qc, Jc = teacher.predict_device(part, batch_size=min(chunk, 256))
        q[lo:hi] = qc + noise * torch.randn(qc.shape, generator=gen, device=dev)
        J[lo:hi] = Jc + (noise / math.sqrt(arch.d_in)) * torch.randn(Jc.shape, generator=gen, device=dev)
    teacher.core.invalidate_plans()
    return DeviceDataset.from_device(_lib.JAC_FULL, m, q, J), tw, m_host


def build_model(arch: Arch, weights=None, name_prefix="", jac="full", precision=None):
    """A DinoModel of ``arch`` (through the reference-named builders) with ``weights`` injected."""
    from .construction.neuralNetworks import projected_dense, projected_resnet, generic_dense
    from .construction.jacobianConstruction import equip_model_with_full_jacobian
    last = None
    if arch.frozen_output and weights is not None:
        last = [np.asarray(w) for w in weights["output_layer"]]
    elif arch.frozen_output:
        last = [np.eye(arch.r_out, arch.d_out, dtype=np.float32), np.zeros(arch.d_out, np.float32)]
    if arch.kind == "generic_dense":
        net = generic_dense(arch.d_in, arch.d_out, list(arch.hidden), name_prefix=name_prefix)
    elif arch.kind == "rb_dense":
        net = projected_dense(last_layer_weights=last, hidden_layer_dimensions=list(arch.hidden),
                              name_prefix=name_prefix, reduced_input_dim=arch.d_in,
                              reduced_output_dim=arch.r_out or arch.d_out)
    else:
        net = projected_resnet(None, last, ranks=list(arch.ranks), name_prefix=name_prefix,
                               reduced_input_dim=arch.d_in, reduced_output_dim=arch.r_out or arch.d_out,
                               last_layer_trainable=False)
    if weights is not None:
        for name, w in weights.items():
            net.get_layer(name_prefix + name).set_weights(w)
    if precision is not None:
        net.precision = precision
    if jac == "full":
        return equip_model_with_full_jacobian(net, name_prefix=name_prefix)
    return net


def flops_per_sample(arch: Arch) -> float:
    """Algorithmic FLOPs of one training sample, SURVEY.md 8d:
    F = (s+1) * 2 * (S + S_minus_first + S_trainable), s = min(d_in, d_out)."""
    table = layer_table(arch)
    sizes = [k * n for _, k, n, *_ in table]
    S = sum(sizes)
    S_t = sum(sz for sz, row in zip(sizes, table) if not (row[0] == "output_layer" and arch.frozen_output))
    s = min(arch.d_in, arch.d_out)
    return (s + 1) * 2.0 * (S + (S - sizes[0]) + S_t)


def chain_flops_per_sample(arch: Arch) -> float:
    """Algorithmic FLOPs of the Jacobian chain alone (forward chain + data-gradient chain):
    s * 2 * (S + S_minus_first), the first two terms of SURVEY.md 8d's F_chain."""
    sizes = [k * n for _, k, n, *_ in layer_table(arch)]
    S = sum(sizes)
    return min(arch.d_in, arch.d_out) * 2.0 * (S + (S - sizes[0]))


def mlp_forward_flops_per_sample(arch: Arch) -> float:
    """Algorithmic FLOPs of the network forward pass of one sample: 2 * S."""
    return 2.0 * sum(k * n for _, k, n, *_ in layer_table(arch))


def bytes_per_sample(arch: Arch) -> float:
    """Compulsory HBM bytes of one sample: read m, q and the target J once (fp32)."""
    return 4.0 * (arch.d_in + arch.d_out + arch.d_out * arch.d_in)
