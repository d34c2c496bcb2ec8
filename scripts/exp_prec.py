"""Prints the actual error levels behind the parity bars: one-step gradient errors and 100-Adam-step parameter errors."""
import sys
import numpy as np, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import helpers as H
from oracle import dino_oracle as O
from dino_b200 import synthetic as S, Adam, SGD
from dino_b200.construction import normalized_mse, normalized_mse_matrix, l2_accuracy, f_accuracy_matrix

def comp(model, opt):
    model.compile(optimizer=opt, loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0], metrics=[l2_accuracy, f_accuracy_matrix])

for cfg, B in (("cfg1", 32), ("cfg2", 256), ("cfg5", 6)):
    arch = S.CONFIGS[cfg]["arch"]
    data, tw = S.make_dataset(arch, B, 40)
    w = S.student_weights(arch, 41, tw, bias_noise=0.1)
    net = H.oracle_net(arch, w)
    qo, Jo = O.batch_jacobian(net, torch.tensor(data["m"], dtype=torch.float64), False)
    (lt, l2, h1), grads = O.h1_step_grads(net, H.oracle_batch(data), (1.0, 1.0), vectorised=True)
    for prec in ("fp32", "fp32x3", "tf32"):
        model = S.build_model(arch, w, precision=prec)
        q, J = model.predict(data["m"], batch_size=B)
        comp(model, SGD(1.0))
        before = {l.name: l.get_weights() for l in model.layers}
        hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=1, batch_size=B, perms=[np.arange(B)])
        gi, errs = 0, []
        for l, lname in enumerate(model.core.names):
            if not model.core.trainable[l]:
                continue
            for a, b in zip(before[lname], model.get_layer(lname).get_weights()):
                errs.append(H.rel_err(a - b, grads[gi].numpy())); gi += 1
        print(f"{cfg} B={B} {prec:7s} J {H.rel_err(J, Jo.numpy()):.2e} loss {abs(hist.history['loss'][0]-lt)/abs(lt):.2e} grads max {max(errs):.2e} median {np.median(errs):.2e}", flush=True)

arch = S.CONFIGS["cfg1"]["arch"]
n = 400
data, tw = S.make_dataset(arch, n, 17)
w = S.student_weights(arch, 18, tw)
perms = [np.random.default_rng(e).permutation(n) for e in range(4)]
net = H.oracle_net(arch, w)
snaps = {}
for steps_epochs in (1, 4):
    net = H.oracle_net(arch, w)
    O.fit(net, H.oracle_batch(data), epochs=steps_epochs, batch_size=16, lr=1e-3, opt="adam", perms=perms, vectorised=True)
    snaps[steps_epochs] = net.get_weights()
for prec in ("fp32", "fp32x3", "tf32"):
    for ep in (1, 4):
        model = S.build_model(arch, w, precision=prec)
        comp(model, Adam(1e-3))
        model.fit([data["m"]], [data["q"], data["J"]], epochs=ep, batch_size=16, perms=perms)
        errs = []
        for layer in model.layers:
            w0 = w.get(layer.name)
            for i, (a, b) in enumerate(zip(layer.get_weights(), snaps[ep].get(layer.name, []))):
                errs.append(H.rel_err(a - w0[i], b - w0[i]))
        print(f"adam {25*ep} steps {prec:7s} update err max {max(errs):.2e} median {np.median(errs):.2e}", flush=True)
