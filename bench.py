#!/usr/bin/env python
"""bench.py - DINO H1 training-step throughput (BASELINE.json metric) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config cfg2]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one optimiser step (forward + per-sample Jacobian + L2/H1 loss + double backward +
[gradient all-reduce] + Adam) on one global minibatch of synthetic teacher data.  Workload at any
N: BASELINE.json configs[1] (confusion-shaped rb_dense, r_in=100, d_Q=50, reduced-input training,
frozen output layer), per-GPU minibatch 256 (weak scaling: global batch 256*N).

Prints ONE JSON line (rank 0).  `value` = samples/s with the dataset resident in HBM, CUDA-event
timed, max over ranks.  `e2e` = the same step through DinoModel.fit_generator with HOST (pinned)
minibatches: every step's H2D copy of (m, q, J) and D2H read of its loss row are inside the timed region
(host clock around the call; the copy of minibatch k+1 overlaps step k, as a tf.data prefetch would);
`e2e.train_on_batch_value` is the strictly serial per-step call for comparison.
`--impl reference` times the reference's CPU path (TensorFlow is not installable here, so the
oracle restatement in fp32 on all host cores, labelled kind="port").
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
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "DINO train samples/s (fwd+Jacobian loss+bwd)"
UNIT = "samples/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg2")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU minibatch (default: the config's)")
    ap.add_argument("--n", type=int, default=16384, help="synthetic samples resident per GPU")
    ap.add_argument("--precision", default=os.environ.get("DINO_B200_PRECISION", "tf32"),
                    help="tf32: tcgen05 tensor-core chain (gradient tolerance 1e-2, tests/test_gpu_tc.py); "
                         "fp32: exact FFMA engine (1e-4 parity mode, tests/test_gpu_parity.py)")
    ap.add_argument("--no-fp32-leg", action="store_true", help="skip the secondary fp32-exact measurement")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--random-data", action="store_true",
                    help="N(0,1) m/q/J instead of teacher data (wide configs: the host-side teacher Jacobian is slow)")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tensor=d["bf16_tflops_sustained"], src="measured (MEASURED_PEAKS.json, bf16 sustained)")
    return dict(hbm=6650.0, tensor=1400.0, src="fallback (B200_PROFILING.md)")


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


def cpu_reference_rate(arch, data, weights, batch, seconds, max_steps=None, threads=None):
    """The reference's CPU training step restated in torch fp32 (oracle/dino_oracle.py: d_out reverse
    passes vectorised like TF's pfor, autograd double backward, Keras Adam).  Returns samples/s."""
    import torch
    import helpers as H
    from oracle import dino_oracle as O
    torch.set_num_threads(threads or os.cpu_count() or 1)
    net = H.oracle_net(arch, weights, dtype=torch.float32)
    opt = O.KerasAdam(net.trainable_tensors(), 1e-3)
    n = data["m"].shape[0]
    full = H.oracle_batch(data, None, dtype=torch.float32)

    def one(i):
        idx = (np.arange(batch) + i * batch) % n
        _, grads = O.h1_step_grads(net, O.slice_batch(full, idx), (1.0, 1.0), vectorised=True)
        opt.step(grads)

    one(0)                                     # warm-up (allocator, thread pool)
    t0, steps = time.perf_counter(), 0
    while True:
        one(steps + 1)
        steps += 1
        el = time.perf_counter() - t0
        if el >= seconds or (max_steps and steps >= max_steps):
            break
    return steps * batch / el, steps, el, torch.get_num_threads()


def run_reference(args, arch, batch):
    """--impl reference: the CPU path on this box's host cores, same config/metric/unit."""
    import warnings
    warnings.filterwarnings("ignore")
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from dino_b200 import synthetic as S
    import helpers as H
    n = max(4 * batch, 1024)
    data, tw = S.make_dataset(arch, n, 1234 + 1)
    w = H.student_weights(arch, 1234 + 2, tw)
    import torch
    from oracle import dino_oracle as O
    # torchrun exports OMP_NUM_THREADS=1; the reference arm is entitled to every host core
    torch.set_num_threads(os.cpu_count() or 1)
    net = H.oracle_net(arch, w, dtype=torch.float32)
    opt = O.KerasAdam(net.trainable_tensors(), 1e-3)
    full = H.oracle_batch(data, None, dtype=torch.float32)

    def one(i):
        idx = (np.arange(batch) + i * batch) % n
        _, grads = O.h1_step_grads(net, O.slice_batch(full, idx), (1.0, 1.0), vectorised=True)
        opt.step(grads)

    # bound the run to a few minutes: cap the number of CPU steps actually executed
    one(0)
    t0 = time.perf_counter(); one(1); per = time.perf_counter() - t0
    warm = min(args.warmup, max(1, int(20.0 / max(per, 1e-6))))
    steps = min(args.steps, max(3, int(120.0 / max(per, 1e-6))))
    for i in range(warm):
        one(i)
    t0 = time.perf_counter()
    for i in range(steps):
        one(i)
    el = time.perf_counter() - t0
    val = steps * batch / el
    cores = torch.get_num_threads()
    sample = f"{steps} steps of {batch} samples, torch-CPU fp32 restatement of the TF path (TF 2.7 not installable)"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": 1e3 * el / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": workload_name(args.config, arch, batch), "per_gpu_batch": batch, "optimizer": "adam"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_name(cfg, arch, batch):
    return (f"{cfg}: {arch.kind} r_in={arch.d_in} d_Q={arch.d_out} hidden={list(arch.hidden) or list(arch.ranks)} "
            f"full reduced-Jacobian H1 loss, per-GPU minibatch {batch}")


def main():
    args = parse()
    from dino_b200 import synthetic as S
    cfg = S.CONFIGS[args.config]
    arch = cfg["arch"]
    batch = args.batch or cfg["batch"]
    if args.impl == "reference":
        run_reference(args, arch, batch)
        return

    import torch
    import helpers as H
    from dino_b200 import parallel, _lib, Adam
    from dino_b200.construction import normalized_mse, normalized_mse_matrix, l2_accuracy, f_accuracy_matrix
    dp = parallel.init_from_env("nccl")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    world = dp.world
    peaks = load_peaks()

    # ---- synthetic teacher data (same on every rank), student initial weights ------------------------
    n = args.n
    if args.random_data:
        rng = np.random.default_rng(1234)
        data = {"m": rng.standard_normal((n, arch.d_in), dtype=np.float32),
                "q": rng.standard_normal((n, arch.d_out), dtype=np.float32),
                "J": rng.standard_normal((n, arch.d_out, arch.d_in), dtype=np.float32)}
        tw = S.random_weights(arch, rng, teacher=True)
    else:
        data, tw = S.make_dataset(arch, n, 1234 + 1)
    w = H.student_weights(arch, 1234 + 2, tw)
    model = S.build_model(arch, w, precision=args.precision)
    model.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
                  metrics=[l2_accuracy, f_accuracy_matrix])
    ds = model.make_dataset([data["m"]], [data["q"], data["J"]])
    Bg = batch * world
    steps_per_epoch = n // Bg
    order = np.random.default_rng(0).permutation(n).astype(np.int32)
    dev = torch.device("cuda", local)
    dataset_bytes = sum(int(t.numel()) * 4 for t in (ds.m, ds.q, ds.J))

    sampler = ClockSampler(local)
    if dp.rank == 0:
        sampler.start()

    eng = model.make_engine(ds, Bg)
    eng.set_order(order)

    def run_steps(k, first):
        for i in range(k):
            if eng.remaining() < Bg:
                eng.rewind()
            eng.step(Bg)

    # ---- resident-data throughput ------------------------------------------------------------------
    run_steps(args.warmup, 0)
    torch.cuda.synchronize(); dp.barrier(); torch.cuda.synchronize()
    sampler.t0 = time.time()
    l0 = _lib.launch_count() + eng.replayed_kernels - eng.captured_kernels
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run_steps(args.steps, args.warmup)
    e1.record()
    torch.cuda.synchronize(); dp.barrier(); torch.cuda.synchronize()
    launches = _lib.launch_count() + eng.replayed_kernels - eng.captured_kernels - l0
    ms = dp.max_over_ranks(e0.elapsed_time(e1))
    ms_per_step = ms / args.steps
    value = Bg * args.steps / (ms * 1e-3)

    # ---- end to end: host minibatches through train_on_batch ------------------------------------------
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    nb_host = min(8, steps_per_epoch)
    host_batches = []
    for s in range(nb_host):
        sel = order[s * Bg + dp.rank * batch: s * Bg + (dp.rank + 1) * batch]
        host_batches.append((pin(data["m"][sel]), pin(data["q"][sel]), pin(data["J"][sel])))
    e2e_steps = args.steps

    def host_stream(k):
        for i in range(k):
            hb = host_batches[i % nb_host]
            yield [hb[0]], [hb[1], hb[2]]

    # (a) one synchronous call per step (Keras train_on_batch: copy, step, read the loss back, return)
    for i in range(max(3, min(args.warmup, 10))):
        hb = host_batches[i % nb_host]
        model.train_on_batch([hb[0]], [hb[1], hb[2]])
    torch.cuda.synchronize(); dp.barrier(); torch.cuda.synchronize()
    k_sync = max(20, e2e_steps // 4)
    t_s0 = time.perf_counter()
    for i in range(k_sync):
        hb = host_batches[i % nb_host]
        model.train_on_batch([hb[0]], [hb[1], hb[2]])
    torch.cuda.synchronize()
    ms_sync = dp.max_over_ranks((time.perf_counter() - t_s0) * 1e3)
    # (b) the streaming call (Keras fit_generator): same copies and read-backs every step, but the copy of minibatch
    #     k+1 overlaps the step on minibatch k.  Timed on the host clock around the whole call, devices idle on both sides.
    model.fit_generator(host_stream(max(3, min(args.warmup, 10))))
    torch.cuda.synchronize(); dp.barrier(); torch.cuda.synchronize()
    t_e0 = time.perf_counter()
    rows = model.fit_generator(host_stream(e2e_steps))
    torch.cuda.synchronize()
    ms_e2e = dp.max_over_ranks((time.perf_counter() - t_e0) * 1e3)
    dp.barrier()
    sampler.t1 = time.time()
    last = rows[-1]
    e2e_value = Bg * e2e_steps / (ms_e2e * 1e-3)
    h2d, d2h = model.last_h2d_bytes, model.last_d2h_bytes
    sampler.stop()

    # ---- dominant kernel, timed alone with CUDA events (eager steps; the timed region above replays a graph) ----
    import ctypes as C
    kern = None
    if args.precision == "tf32" and world == 1:
        prof_eng = model.make_engine(ds, Bg, use_graph=False)
        prof_eng.set_order(order)
        lib = _lib.load()
        ms_sum, cnt = C.c_double(0.0), C.c_int64(0)
        for i in range(5):
            if prof_eng.remaining() < Bg:
                prof_eng.rewind()
            prof_eng.step(Bg)
        _lib.check(lib.dino_plan_profile(prof_eng.plan.handle, 1, C.byref(ms_sum), C.byref(cnt)), "dino_plan_profile")
        for i in range(min(args.steps, 200)):
            if prof_eng.remaining() < Bg:
                prof_eng.rewind()
            prof_eng.step(Bg)
        torch.cuda.synchronize()
        _lib.check(lib.dino_plan_profile(prof_eng.plan.handle, 0, C.byref(ms_sum), C.byref(cnt)), "dino_plan_profile")
        if cnt.value > 0:
            kern = {"name": "chain_fused_kernel", "ms": ms_sum.value / cnt.value, "launches_timed": int(cnt.value)}

    # ---- secondary leg: the exact-fp32 engine on the same workload (1e-4 parity mode) ----
    fp32_leg = None
    if args.precision != "fp32" and not args.no_fp32_leg:
        m32 = S.build_model(arch, w, precision="fp32")
        m32.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
                    metrics=[l2_accuracy, f_accuracy_matrix])
        e32 = m32.make_engine(ds, Bg)
        e32.set_order(order)
        k32 = max(20, args.steps // 4)
        for i in range(max(3, args.warmup // 2)):
            if e32.remaining() < Bg:
                e32.rewind()
            e32.step(Bg)
        torch.cuda.synchronize(); dp.barrier(); torch.cuda.synchronize()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for i in range(k32):
            if e32.remaining() < Bg:
                e32.rewind()
            e32.step(Bg)
        f1.record()
        torch.cuda.synchronize(); dp.barrier(); torch.cuda.synchronize()
        ms32 = dp.max_over_ranks(f0.elapsed_time(f1))
        fp32_leg = {"value": Bg * k32 / (ms32 * 1e-3), "unit": UNIT, "ms_per_step": ms32 / k32, "steps": k32,
                    "dtype": "fp32", "tolerance": "1e-4 relative vs the fp64 oracle (tests/test_gpu_parity.py)"}

    if dp.rank != 0:
        return
    # ---- roofline of the step (all kernels of one step taken as the unit) ----------------
    F = S.flops_per_sample(arch)
    achieved_tf = F * batch / (ms_per_step * 1e-3) / 1e12          # per GPU
    step_roof = {"achieved_tflops": achieved_tf, "frac_of_peak": achieved_tf / peaks["tensor"], "flops_per_sample": F,
                 "hbm_achieved_gbs": S.bytes_per_sample(arch) * batch / (ms_per_step * 1e-3) / 1e9,
                 "hbm_peak_gbs": peaks["hbm"], "note": "whole step (all launches), algorithmic FLOPs of SURVEY 8d"}
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        ent = tj.get("chain_fused_kernel")
        if ent and ent.get("per_gpu_batch") == batch and ent.get("config") == args.config:
            traffic = ent["dram_bytes_per_launch"]
    if kern is not None:
        # forward chain + data-gradient chain (SURVEY 8d terms) + the network forward pass the kernel also runs
        Fk = S.chain_flops_per_sample(arch) + S.mlp_forward_flops_per_sample(arch)
        ach = Fk * batch / (kern["ms"] * 1e-3) / 1e12
        roof = {"bound": "tensor", "achieved": ach, "peak": peaks["tensor"], "unit": "TFLOP/s",
                "frac": ach / peaks["tensor"], "traffic": traffic, "peak_source": peaks["src"],
                "kernel": "chain_fused_kernel (network forward + output loss, tcgen05 Jacobian chain: forward + H1 loss + data-gradient backward)",
                "kernel_ms": kern["ms"], "launches_timed": kern["launches_timed"],
                "algorithmic_flops_per_launch": Fk * batch,
                "timing": "CUDA events on the launching stream around the kernel, eager steps after the timed region",
                "tf32_cublas_peak_measured_tflops": 713.0, "step": step_roof}
    else:
        roof = {"bound": "tensor", "achieved": achieved_tf, "peak": peaks["tensor"], "unit": "TFLOP/s",
                "frac": achieved_tf / peaks["tensor"], "traffic": None, "peak_source": peaks["src"],
                "kernel": "whole H1 step (all launches of one step), algorithmic FLOPs of SURVEY 8d", "step": step_roof}
    # ---- the chain contraction on its own at a wide layer (cfg5's 1024-wide hidden layers, 256 samples x s = 100 rows):
    #      the tcgen05 engine next to cuBLAS TF32 on the same operands, CUDA-event timed.  north_star's ">= 50 % of dense
    #      tensor-core peak on the Jacobian contraction" is about this kernel; cfg2's 50-100 wide layers cannot get there.
    if args.precision == "tf32" and world == 1 and not args.no_fp32_leg:
        Mg, Ng, Kg = 25600, 1024, 1024
        Ag = torch.randn(Mg, Kg, device=dev); Bg_ = torch.randn(Ng, Kg, device=dev); Cg = torch.empty(Mg, Ng, device=dev)

        def _time(fn, iters=20):
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

        from dino_b200.model import _stream
        ms_ours = _time(lambda: _lib.check(lib.dino_gemm_nt(_lib.PREC_TF32, C.c_void_p(Ag.data_ptr()), Kg, C.c_void_p(Bg_.data_ptr()),
                                                            Kg, C.c_void_p(Cg.data_ptr()), Ng, Mg, Ng, Kg, _stream()), "gemm"))
        old_tf32 = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = True
        ms_cublas = _time(lambda: torch.matmul(Ag, Bg_.t()))
        torch.backends.cuda.matmul.allow_tf32 = old_tf32
        flg = 2.0 * Mg * Ng * Kg
        roof["contraction_engine"] = {"kernel": "gemm_tc_kernel (TMA -> tcgen05.mma kind::tf32 -> TMEM), dino_gemm_nt",
                                      "shape_MNK": [Mg, Ng, Kg], "achieved_tflops": flg / ms_ours / 1e9,
                                      "cublas_tf32_tflops_same_operands": flg / ms_cublas / 1e9,
                                      "frac_of_cublas_tf32": ms_cublas / ms_ours,
                                      "frac_of_bf16_peak": flg / ms_ours / 1e9 / peaks["tensor"]}
        del Ag, Bg_, Cg
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        import warnings
        warnings.filterwarnings("ignore")
        sub = {k: v[:max(4 * batch, 1024)] for k, v in data.items()}
        rate, csteps, el, cores = cpu_reference_rate(arch, sub, w, batch, args.cpu_seconds)
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{csteps} steps of {batch} samples in {el:.1f} s: torch-CPU fp32 restatement of the "
                         f"reference's TF step (reverse-mode batch Jacobian, autograd double backward, Keras Adam)"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.precision,
            "data": "synthetic",
            "config": {"workload": workload_name(args.config, arch, batch), "per_gpu_batch": batch,
                       "global_batch": Bg, "resident_samples": n, "optimizer": "adam", "precision": args.precision,
                       "parallelism": f"dp{world}", "cuda_graph_step": bool(eng.use_graph and eng.graphs),
                       "tolerance": ("gradients within 1e-2, Jacobian within 5e-3 of the fp64 oracle (tests/test_gpu_tc.py)"
                                     if args.precision == "tf32" else "1e-4 relative vs the fp64 oracle (tests/test_gpu_parity.py)"),
                       "l2_policy": f"inputs larger than L2: {dataset_bytes / 1e6:.0f} MB dataset cycled, "
                                    f"a different minibatch every step"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / e2e_steps, "loss_last": last[0],
                    "api": "DinoModel.fit_generator over pinned host minibatches (copy of batch k+1 overlaps step k; every "
                           "step's loss row is read back)",
                    "train_on_batch_value": Bg * k_sync / (ms_sync * 1e-3), "timing": "host clock around the call"},
            "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu, "fp32_exact": fp32_leg,
            "clocks": sampler.summary()}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
