#!/usr/bin/env python
"""bench.py - DINO H1 training-step throughput (BASELINE.json metric) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config cfg5] [--precision fp32x3]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one optimiser step (forward + per-sample Jacobian + L2/H1 loss + double backward +
[gradient all-reduce] + Adam) on one global minibatch of synthetic teacher data.  Workload at every N:
BASELINE.json configs[4] (scaling sweep: rb_dense r_in=200, r_out=100, hidden width 1024, 64k synthetic
samples), per-GPU minibatch 256 (weak scaling: global batch 256*N).  Arithmetic: fp32x3 (tcgen05, three
TF32 MMAs on hi/lo splits of both operands; tested against the fp64 oracle at 1e-4).

Prints ONE JSON line (rank 0).  `value` = samples/s with the dataset resident in HBM, CUDA-event
timed, max over ranks.  `e2e` = the same step through DinoModel.fit_generator with HOST (pinned)
minibatches: every step's H2D copy of (m, q, J) and D2H read of its loss row are inside the timed region.
Secondary legs (one GPU only): `extra.tf32` (same workload, single-pass TF32), `legs.cfg2` (BASELINE
configs[1], the fused chain kernel), `legs.cfg4` (configs[3], full Jacobian, HBM-bound), the contraction
engine next to cuBLAS TF32, and the CPU baseline (oracle port on the host cores, bounded sample).
`--impl reference` times the reference's CPU path (TensorFlow is not installable here, so the oracle
restatement in fp32 on all host cores, labelled kind="port").
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "DINO train samples/s (fwd+Jacobian loss+bwd)"
UNIT = "samples/s"
TOLERANCE = {"fp32x3": "Jacobian, loss and every gradient within 1e-4 of the fp64 oracle at every BASELINE shape (tests/test_gpu_shapes.py)",
             "fp32": "1e-4 relative vs the fp64 oracle (tests/test_gpu_parity.py, tests/test_gpu_shapes.py)",
             "tf32": "Jacobian within 5e-3, gradients within 1e-2 of the fp64 oracle (tests/test_gpu_shapes.py)"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg5")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU minibatch (default: the config's)")
    ap.add_argument("--n", type=int, default=0, help="synthetic samples resident per GPU (default: the config's)")
    ap.add_argument("--precision", default=os.environ.get("DINO_B200_PRECISION", "fp32x3"),
                    help="fp32x3 (default; 1e-4 parity mode on the tensor cores), tf32 (reported variant), fp32 (SIMT)")
    ap.add_argument("--no-legs", action="store_true", help="skip the secondary legs (tf32, cfg2, cfg4, contraction engine)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops_sustained"], bf16_burst=d.get("bf16_tflops", d["bf16_tflops_sustained"]),
                    src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16=1400.0, bf16_burst=1590.0, src="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed regions run."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index
        self.t0 = self.t1 = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()

    def summary(self):
        rows = [r for t, r in self.rows if self.t0 is not None and self.t0 - 0.05 <= t <= (self.t1 or 1e30) + 0.25]
        if not rows:
            rows = [r for _, r in self.rows]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(float(r[1]) for r in rows)
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            for nm, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][2]), "reasons": sorted(reasons),
                "samples": len(rows), "power_w_max": max(float(r[3]) for r in rows)}


def workload_name(cfg, arch, batch):
    jac = "full Jacobian (unreduced input)" if arch.kind == "generic_dense" else "full reduced-Jacobian"
    return (f"{cfg}: {arch.kind} d_in={arch.d_in} d_Q={arch.d_out} hidden={list(arch.hidden) or list(arch.ranks)} "
            f"{jac} H1 loss, per-GPU minibatch {batch}")


def config_dict(cfg, arch, batch, world, n):
    """The `config` object of the JSON line: identical in both arms (--impl ours / reference) for the same arguments."""
    from dino_b200 import synthetic as S
    mb = n * S.bytes_per_sample(arch) / 1e6
    return {"workload": workload_name(cfg, arch, batch), "per_gpu_batch": batch, "global_batch": batch * world,
            "resident_samples": n, "optimizer": "adam", "parallelism": f"dp{world}",
            "l2_policy": f"inputs larger than L2: {mb:.0f} MB dataset cycled, a different minibatch every step"}


def cpu_batch_for(arch, batch):
    """Samples per CPU step of the bounded reference sample: the full minibatch for the narrow configurations, fewer
    for the wide ones (a 256-sample step of the width-1024 network takes the host cores tens of seconds)."""
    from dino_b200 import synthetic as S
    return batch if S.flops_per_sample(arch) * batch < 5e10 else max(8, min(batch, int(5e10 / S.flops_per_sample(arch))))


def _cpu_problem(arch, n, seed=1235):
    """Host-side data for the CPU arm: teacher data for the narrow configurations; same-shape N(0,1) arrays for the wide
    ones (the numpy teacher Jacobian of a width-1024 network costs minutes; the step's cost does not depend on values)."""
    from dino_b200 import synthetic as S
    if S.flops_per_sample(arch) < 1e8:
        data, tw = S.make_dataset(arch, n, seed)
        return data, S.student_weights(arch, seed + 1, tw), "teacher data"
    rng = np.random.default_rng(seed)
    data = {"m": rng.standard_normal((n, arch.d_in), dtype=np.float32),
            "q": rng.standard_normal((n, arch.d_out), dtype=np.float32),
            "J": rng.standard_normal((n, arch.d_out, arch.d_in), dtype=np.float32)}
    tw = S.random_weights(arch, rng, teacher=True)
    return data, S.student_weights(arch, seed + 1, tw), "N(0,1) data of the same shapes"


def _cpu_stepper(arch, data, weights, batch):
    """One training step of the reference's CPU path restated in torch fp32 (oracle/dino_oracle.py: d_out reverse
    passes vectorised like TF's pfor, autograd double backward, Keras Adam) on all host cores."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H
    from oracle import dino_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)      # (torchrun exports OMP_NUM_THREADS=1; this arm is entitled to every core)
    net = H.oracle_net(arch, weights, dtype=torch.float32)
    opt = O.KerasAdam(net.trainable_tensors(), 1e-3)
    n = data["m"].shape[0]
    full = H.oracle_batch(data, None, dtype=torch.float32)

    def one(i):
        idx = (np.arange(batch) + i * batch) % n
        _, grads = O.h1_step_grads(net, O.slice_batch(full, idx), (1.0, 1.0), vectorised=True)
        opt.step(grads)
    return one, torch.get_num_threads()


def cpu_reference_rate(arch, batch, seconds):
    """cpu_baseline of the product arm: a bounded sample (about `seconds` of CPU work) of the same workload."""
    import warnings
    warnings.filterwarnings("ignore")
    cb = cpu_batch_for(arch, batch)
    data, w, what = _cpu_problem(arch, max(4 * cb, 64))
    one, cores = _cpu_stepper(arch, data, w, cb)
    one(0)                                     # warm-up (allocator, thread pool)
    t0, steps = time.perf_counter(), 0
    while True:
        one(steps + 1)
        steps += 1
        el = time.perf_counter() - t0
        if el >= seconds:
            break
    return {"value": steps * cb / el, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{steps} steps of {cb} samples in {el:.1f} s ({what}): torch-CPU fp32 restatement of the reference's "
                      f"TF step (reverse-mode batch Jacobian, autograd double backward, Keras Adam); TF 2.7 is not installable here"}


def run_reference(args, cfg_name, arch, batch, n):
    """--impl reference: the CPU path on this box's host cores, same config / metric / unit; rank 0 only."""
    import warnings
    warnings.filterwarnings("ignore")
    if int(os.environ.get("RANK", "0")) != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cb = cpu_batch_for(arch, batch)
    data, w, what = _cpu_problem(arch, max(4 * cb, 64))
    one, cores = _cpu_stepper(arch, data, w, cb)
    # bound the run to a few minutes: cap the number of CPU steps actually executed
    one(0)
    t0 = time.perf_counter(); one(1); per = time.perf_counter() - t0
    warm = min(args.warmup, max(1, int(15.0 / max(per, 1e-6))))
    steps = min(args.steps, max(3, int(90.0 / max(per, 1e-6))))
    for i in range(warm):
        one(i)
    t0 = time.perf_counter()
    for i in range(steps):
        one(i)
    el = time.perf_counter() - t0
    val = steps * cb / el
    sample = (f"{steps} steps of {cb} samples ({what}), torch-CPU fp32 restatement of the TF path on {cores} host threads "
              f"(TF 2.7 not installable)")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": 1e3 * el / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": config_dict(cfg_name, arch, batch, world, n),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------------------
# GPU measurements
# --------------------------------------------------------------------------------------------------------------
def _compile(model):
    from dino_b200 import Adam
    from dino_b200.construction import normalized_mse, normalized_mse_matrix, l2_accuracy, f_accuracy_matrix
    model.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
                  metrics=[l2_accuracy, f_accuracy_matrix])


def resident_steps(model, ds, Bg, steps, warmup, dp, order):
    """K optimiser steps over the resident dataset (device-side cursors, CUDA-graph replay), CUDA-event timed."""
    import torch
    from dino_b200 import _lib
    eng = model.make_engine(ds, Bg)
    eng.set_order(order)

    def run(k):
        for _ in range(k):
            if eng.remaining() < Bg:
                eng.rewind()
            eng.step(Bg)

    run(max(warmup, 3))
    torch.cuda.synchronize(); dp.barrier(); torch.cuda.synchronize()
    l0 = _lib.launch_count() + eng.replayed_kernels - eng.captured_kernels
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(steps)
    e1.record()
    torch.cuda.synchronize(); dp.barrier(); torch.cuda.synchronize()
    launches = _lib.launch_count() + eng.replayed_kernels - eng.captured_kernels - l0
    ms = dp.max_over_ranks(e0.elapsed_time(e1))
    return {"ms_per_step": ms / steps, "value": Bg * steps / (ms * 1e-3), "launches": int(launches),
            "graph": bool(eng.use_graph and eng.graphs)}


def e2e_steps(model, ds, batch, Bg, steps, warmup, dp, order):
    """The same step through the public API with HOST minibatches (pinned): DinoModel.fit_generator copies minibatch k+1
    host->device while step k runs and reads every step's loss row back; host clock around the whole call."""
    import torch
    nb_host = max(2, min(8, ds.n // Bg))
    host = []
    for s in range(nb_host):
        sel = torch.from_numpy(order[s * Bg + dp.rank * batch: s * Bg + (dp.rank + 1) * batch].astype(np.int64)).to(ds.m.device)
        host.append(tuple(t.index_select(0, sel).cpu().pin_memory() for t in (ds.m, ds.q, ds.J)))

    def stream(k):
        for i in range(k):
            hb = host[i % nb_host]
            yield [hb[0]], [hb[1], hb[2]]

    model.fit_generator(stream(max(3, min(warmup, 10))))
    torch.cuda.synchronize(); dp.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    rows = model.fit_generator(stream(steps))
    torch.cuda.synchronize()
    ms = dp.max_over_ranks((time.perf_counter() - t0) * 1e3)
    dp.barrier()
    return {"value": Bg * steps / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": model.last_h2d_bytes,
            "d2h_bytes_per_step": model.last_d2h_bytes, "ms_per_step": ms / steps, "loss_last": rows[-1][0],
            "api": "DinoModel.fit_generator over pinned host minibatches (copy of batch k+1 overlaps step k; every step's "
                   "loss row is read back)", "timing": "host clock around the call"}


def chain_stage_ms(model, ds, Bg, steps, order):
    """CUDA events on the launching stream around the Jacobian-chain stage of eager steps (dino_plan_profile)."""
    import ctypes as C
    import torch
    from dino_b200 import _lib
    eng = model.make_engine(ds, Bg, use_graph=False)
    eng.set_order(order)
    lib = _lib.load()
    ms_sum, cnt = C.c_double(0.0), C.c_int64(0)

    def run(k):
        for _ in range(k):
            if eng.remaining() < Bg:
                eng.rewind()
            eng.step(Bg)

    run(3)
    _lib.check(lib.dino_plan_profile(eng.plan.handle, 1, C.byref(ms_sum), C.byref(cnt)), "dino_plan_profile")
    run(steps)
    torch.cuda.synchronize()
    _lib.check(lib.dino_plan_profile(eng.plan.handle, 0, C.byref(ms_sum), C.byref(cnt)), "dino_plan_profile")
    return (ms_sum.value / cnt.value, int(cnt.value)) if cnt.value > 0 else (None, 0)


def time_ms(fn, iters=20):
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for _ in range(iters):
        fn()
    g1.record()
    torch.cuda.synchronize()
    return g0.elapsed_time(g1) / iters


def contraction_engine(dev):
    """The chain contraction on its own at cfg5's layer shape (256 samples x s = 100 rows, width 1024), both arithmetic
    modes, next to cuBLAS TF32 (torch.matmul) on the same operands: the measured TF32 peak of this GPU."""
    import ctypes as C
    import torch
    from dino_b200 import _lib
    from dino_b200.model import _stream
    lib = _lib.load()
    M, N, K = 25600, 1024, 1024
    A = torch.randn(M, K, device=dev); Bt = torch.randn(N, K, device=dev); Cm = torch.empty(M, N, device=dev)
    fl = 2.0 * M * N * K
    out = {"kernel": "gemm_tc_kernel (TMA -> [hi/lo split] -> tcgen05.mma kind::tf32 -> TMEM), dino_gemm_nt", "shape_MNK": [M, N, K]}
    for name, eng in (("tf32", _lib.PREC_TF32), ("fp32x3", _lib.PREC_FP32X3)):
        ms = time_ms(lambda: _lib.check(lib.dino_gemm_nt(eng, C.c_void_p(A.data_ptr()), K, C.c_void_p(Bt.data_ptr()), K,
                                                         C.c_void_p(Cm.data_ptr()), N, M, N, K, _stream()), "gemm"))
        out[name + "_tflops"] = fl / ms / 1e9
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    ms = time_ms(lambda: torch.matmul(A, Bt.t()))
    torch.backends.cuda.matmul.allow_tf32 = old
    out["cublas_tf32_tflops_same_operands"] = fl / ms / 1e9
    out["tf32_frac_of_cublas_tf32"] = out["tf32_tflops"] / out["cublas_tf32_tflops_same_operands"]
    out["fp32x3_frac_of_cublas_tf32_over_3"] = 3.0 * out["fp32x3_tflops"] / out["cublas_tf32_tflops_same_operands"]
    return out


def n_chain_launches(arch):
    """Kernels between the two profile events of the layer-by-layer path: one contraction per layer forward, the H1 loss
    kernel (+ the output-loss kernel), one contraction per layer backward (none for a linear top layer)."""
    from dino_b200 import synthetic as S
    L = len(S.layer_table(arch))
    return L + 2 + (L - 1)


def build_problem(cfg_name, n, precision, ds=None, tw=None):
    from dino_b200 import synthetic as S
    arch = S.CONFIGS[cfg_name]["arch"]
    if ds is None:
        ds, tw, _ = S.make_device_dataset(arch, n, 1235)
    w = S.student_weights(arch, 1236, tw)
    model = S.build_model(arch, w, precision=precision)
    _compile(model)
    return arch, model, ds, tw


def main():
    args = parse()
    from dino_b200 import synthetic as S
    cfg = S.CONFIGS[args.config]
    arch = cfg["arch"]
    batch = args.batch or cfg["batch"]
    n = args.n or cfg["n"]
    if args.impl == "reference":
        run_reference(args, args.config, arch, batch, n)
        return

    import torch
    from dino_b200 import parallel
    dp = parallel.init_from_env("nccl")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    world = dp.world
    peaks = load_peaks()
    Bg = batch * world
    if n < 2 * Bg:
        n = 2 * Bg
    order = np.random.default_rng(0).permutation(n).astype(np.int32)

    # ---- synthetic teacher data (same on every rank), produced on the device by this library's own kernels ----
    arch, model, ds, tw = build_problem(args.config, n, args.precision)
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if dp.rank == 0:
        sampler.start()
    sampler.t0 = time.time()
    main_run = resident_steps(model, ds, Bg, args.steps, args.warmup, dp, order)
    e2e = e2e_steps(model, ds, batch, Bg, args.steps, args.warmup, dp, order)
    sampler.t1 = time.time()
    sampler.stop()
    stage_ms, stage_n = chain_stage_ms(model, ds, Bg, min(args.steps, 20), order)
    if dp.rank != 0:
        return                                  # (what follows runs on rank 0 alone, without collectives)

    legs, extra, engine = {}, {}, None
    if not args.no_legs:
        engine = contraction_engine(dev)
    if not args.no_legs and world == 1:
        # the same workload in single-pass TF32 (reported variant, looser tolerance)
        if args.precision != "tf32":
            _, mtf, _, _ = build_problem(args.config, n, "tf32", ds, tw)
            r = resident_steps(mtf, ds, Bg, max(10, args.steps // 2), args.warmup, dp, order)
            extra["tf32"] = {"value": r["value"], "unit": UNIT, "ms_per_step": r["ms_per_step"], "dtype": "tf32",
                             "achieved_tflops": S.flops_per_sample(arch) * batch / (r["ms_per_step"] * 1e-3) / 1e12,
                             "tolerance": TOLERANCE["tf32"]}
            del mtf
        del model, ds
        torch.cuda.empty_cache()
        for leg, ln, lsteps, lb in (("cfg2", 16384, 200, 256), ("cfg4", 512, 12, 128)):
            if leg == args.config:
                continue
            la, lm, lds, ltw = build_problem(leg, ln, args.precision)
            lorder = np.random.default_rng(0).permutation(ln).astype(np.int32)
            r = resident_steps(lm, lds, lb, lsteps, 10, dp, lorder)
            le = e2e_steps(lm, lds, lb, lb, lsteps, 10, dp, lorder)
            ent = {"config": config_dict(leg, la, lb, 1, ln), "dtype": args.precision, "value": r["value"], "unit": UNIT,
                   "ms_per_step": r["ms_per_step"], "gpu_launches_per_step": r["launches"] / lsteps, "e2e": le,
                   "achieved_tflops": S.flops_per_sample(la) * lb / (r["ms_per_step"] * 1e-3) / 1e12,
                   "achieved_hbm_gbs": S.bytes_per_sample(la) * lb / (r["ms_per_step"] * 1e-3) / 1e9}
            sm, sn = chain_stage_ms(lm, lds, lb, min(lsteps, 50), lorder)
            if leg == "cfg2" and sm is not None:
                Fk = S.chain_flops_per_sample(la) + S.mlp_forward_flops_per_sample(la)
                ent["kernel"] = {"name": "chain_fused_kernel", "ms": sm, "launches_timed": sn,
                                 "achieved_tflops": Fk * lb / (sm * 1e-3) / 1e12}
            if leg == "cfg4":
                # HBM-bound configuration: algorithmic bytes = m, q and the target Jacobian read once
                ent["roofline"] = {"bound": "hbm", "achieved": ent["achieved_hbm_gbs"], "peak": peaks["hbm"], "unit": "GB/s",
                                   "frac": ent["achieved_hbm_gbs"] / peaks["hbm"], "traffic": None,
                                   "note": "whole step; floor = 6.78 MB per sample / measured copy bandwidth"}
            if args.precision != "tf32":
                _, ltf, _, _ = build_problem(leg, ln, "tf32", lds, ltw)
                rt = resident_steps(ltf, lds, lb, lsteps, 10, dp, lorder)
                ent["tf32"] = {"value": rt["value"], "ms_per_step": rt["ms_per_step"]}
                del ltf
            legs[leg] = ent
            del lm, lds
            torch.cuda.empty_cache()

    # ---- roofline of the dominant kernel: the chain contractions (gemm_tc_kernel) of the step ---------------------
    F = S.flops_per_sample(arch)
    achieved_tf = F * batch / (main_run["ms_per_step"] * 1e-3) / 1e12          # per GPU
    tf32_peak = engine["cublas_tf32_tflops_same_operands"] if engine else 690.0
    tf32_src = ("cuBLAS TF32 measured live on the same operands (roofline.contraction_engine)" if engine
                else "cuBLAS TF32 at 25600x1024x1024 measured earlier on this pool (profiles/r02b_summary.md)")
    div = 3.0 if args.precision == "fp32x3" else 1.0
    peak = tf32_peak / div
    step_roof = {"achieved_tflops": achieved_tf, "frac_of_peak": achieved_tf / peak, "frac_of_bf16_sustained": achieved_tf / peaks["bf16"],
                 "flops_per_sample": F, "hbm_achieved_gbs": S.bytes_per_sample(arch) * batch / (main_run["ms_per_step"] * 1e-3) / 1e9,
                 "hbm_peak_gbs": peaks["hbm"], "note": "whole step (all launches), algorithmic FLOPs of SURVEY 8d"}
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            ent = json.load(f).get(f"chain_stage:{args.config}:{args.precision}:{batch}")
        if ent:
            traffic = ent["dram_bytes_per_launch"]
    fused = arch.d_in <= 128 and max(list(arch.hidden) + [arch.d_out, arch.d_in]) <= 128
    if stage_ms is not None:
        nl = 1 if fused else n_chain_launches(arch)
        Fk = S.chain_flops_per_sample(arch) + (S.mlp_forward_flops_per_sample(arch) if fused else 0.0)
        ach = Fk * batch / (stage_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic,
                "peak_source": tf32_src + ("; fp32x3 issues three TF32 MMAs per product: peak = TF32 / 3" if div == 3.0 else ""),
                "frac_of_bf16_sustained": ach / peaks["bf16"], "bf16_peak_source": peaks["src"],
                "kernel": ("chain_fused_kernel" if fused else
                           f"gemm_tc_kernel: the {nl - 2} chain contractions of a step (forward chain + data-gradient chain; the two "
                           f"loss kernels between them are inside the timed stage)"),
                "kernel_ms": stage_ms / nl, "stage_ms": stage_ms, "launches_per_stage": nl, "stages_timed": stage_n,
                "algorithmic_flops_per_stage": Fk * batch,
                "timing": "CUDA events on the launching stream around the stage, eager steps after the timed region",
                "step": step_roof}
    else:
        roof = {"bound": "tensor", "achieved": achieved_tf, "peak": peak, "unit": "TFLOP/s", "frac": achieved_tf / peak,
                "traffic": None, "peak_source": tf32_src, "kernel": "whole H1 step (all launches)", "step": step_roof}
    if engine:
        roof["contraction_engine"] = engine
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cpu = cpu_reference_rate(arch, batch, args.cpu_seconds)
    line = {"metric": METRIC, "value": main_run["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": main_run["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": config_dict(args.config, arch, batch, world, n),
            "arithmetic": {"mode": args.precision, "tolerance": TOLERANCE[args.precision], "cuda_graph_step": main_run["graph"]},
            "e2e": e2e, "gpu_launches": main_run["launches"], "roofline": roof, "cpu_baseline": cpu, "extra": extra,
            "legs": legs, "clocks": sampler.summary()}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
