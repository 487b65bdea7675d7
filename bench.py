#!/usr/bin/env python
"""bench.py — training characters/sec of the char-RNN step on N B200s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c3] [--impl reference]

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path over one batch:
embedding gather -> L recurrent layers -> projection/softmax/xent -> BPTT -> (all-reduce)
-> clip + Adam, i.e. the reference's `sess.run([cost, final_state, train_op])`, train.py:199.

  value     chars/s with the id matrices already resident in HBM (device-timed, CUDA events)
  e2e       the same through Model.step() with HOST id buffers: pinned H2D of x,y and a D2H
            read of the loss every step (what train.py's loop does)
  roofline  the dominant phase of the step, timed live with CUDA events (crnn_profile_*)
  cpu_baseline / --impl reference
            the CPU restatement of the reference graph (oracle/torch_restatement.py; TF 1.8
            itself cannot run here) on the box's host cores
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# BASELINE.json configs.  c3 is the default: it is the config the metric
# "train chars/sec at 1/2/4/8 B200" is quoted on ("... at 1/2/4/8 B200"), and it fits one GPU.
CONFIGS = {
    "c1": dict(model="lstm", rnn_size=128, num_layers=2, seq_length=50, batch_size=50, vocab_size=65,
               desc="LSTM 2x128, S=50, B=50/GPU, V=65 (train.py defaults; tinyshakespeare shape)"),
    "c3": dict(model="gru", rnn_size=512, num_layers=2, seq_length=100, batch_size=256, vocab_size=96,
               desc="GRU 2x512, S=100, B=256/GPU, V=96, synthetic Zipf text"),
    "c4": dict(model="lstm", rnn_size=1024, num_layers=3, seq_length=200, batch_size=128, vocab_size=96,
               desc="LSTM 3x1024, S=200, B=128/GPU, V=96, synthetic Zipf text"),
    "c5": dict(model="rnn", rnn_size=2048, num_layers=4, seq_length=256, batch_size=512, vocab_size=256,
               desc="tanh RNN 4x2048, S=256, B=512/GPU, V=256, synthetic text"),
}
GATES = {"lstm": 4, "gru": 3, "rnn": 1}
L2_BYTES = 126e6


def train_flop_per_char(c):
    H, L, V, G = c["rnn_size"], c["num_layers"], c["vocab_size"], GATES[c["model"]]
    return 3 * (L * 2 * (2 * H) * (G * H) + 2 * H * V)


def phase_flops(c):
    """Algorithmic FLOPs per step of each GEMM-shaped phase."""
    H, L, V, G = c["rnn_size"], c["num_layers"], c["vocab_size"], GATES[c["model"]]
    SB = c["seq_length"] * c["batch_size"]
    unit = 2.0 * SB * H * G * H * L
    return {"xproj_gemm": unit, "rec_fwd": unit, "rec_bwd": unit, "input_grads": unit, "weight_grads": 2 * unit,
            "logits_gemm": 2.0 * SB * H * V, "proj_grads": 4.0 * SB * H * V}


def phase_bytes(c, P):
    """Algorithmic HBM bytes per step of the bandwidth-bound phases."""
    H, V = c["rnn_size"], c["vocab_size"]
    SB = c["seq_length"] * c["batch_size"]
    return {"embed_gather": SB * 4 + V * H * 4 + SB * H * 4, "embed_scatter": SB * 4 + SB * H * 4 + V * H * 4,
            "clip_adam": 32.0 * P, "softmax_xent": 2.0 * SB * V * 4 + SB * 4}


def synthetic_ids(c, nb, seed):
    """Zipf(s=1) ids over a seeded permutation of the vocab, as the loader's [B, nb*S] matrices."""
    V, B, S = c["vocab_size"], c["batch_size"], c["seq_length"]
    rng = np.random.Generator(np.random.PCG64(seed))
    w = 1.0 / np.arange(1, V + 1)
    p = w / w.sum()
    ids = rng.permutation(V)[rng.choice(V, size=B * nb * S, p=p)].astype(np.int32)
    X = ids.reshape(B, nb * S)
    Y = np.roll(ids, -1).reshape(B, nb * S)
    return X, Y


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sus=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sus=1400.0, src="fallback")


class ClockSampler(threading.Thread):
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([f.strip() for f in out.strip().split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def summary(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_plan(c, per_step_budget, probe_steps=3):
    """Batch size and thread count of the CPU restatement on this box - ONE decision for both CPU legs
    (``--impl reference`` and ``cpu_baseline``), so they cannot disagree.

    The per-step sample is the full batch unless a full step would exceed ``per_step_budget`` seconds (configs 4/5 are
    tens of TFLOP per step on a CPU), in which case it is halved until it fits.  Thread candidates {all, all/2} are
    each timed over ``probe_steps`` steps (after one warm-up step) AT THAT BATCH; the faster one wins.  The probe is
    returned so the JSON line can show it (``host_threads_probe``)."""
    ncpu = os.cpu_count() or 1
    B = c["batch_size"]
    pb = max(1, B // 8)
    quick = cpu_restatement_rate(c, 1e9, 1, steps=1, batch=pb, threads=ncpu)
    per_row = quick["seconds"] / pb
    bs = B
    while bs > 1 and per_row * bs > per_step_budget:
        bs //= 2
    probe = {}
    for th in sorted({ncpu, max(1, ncpu // 2)}, reverse=True):
        r = cpu_restatement_rate(c, 1e9, 1, steps=probe_steps, batch=bs, threads=th)
        probe[str(th)] = round(r["value"], 1)
    best = max(probe, key=lambda k: probe[k])
    return {"threads": int(best), "batch": bs, "probe_chars_per_s": probe, "probe_steps": probe_steps}


def cpu_restatement_rate(c, max_seconds, warmup, steps=None, batch=None, seed=0, threads=None, data_seed=1234):
    """chars/s of the CPU restatement (oracle/torch_restatement.py) on this box's host cores."""
    import torch
    from oracle import char_rnn_oracle as O
    from oracle.torch_restatement import TorchStep
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    B = batch or c["batch_size"]
    cc = dict(c, batch_size=B)
    cfg = O.Config(c["model"], c["rnn_size"], c["num_layers"], c["seq_length"], B, c["vocab_size"], 5.0)
    ts = TorchStep(cfg, O.init_params(cfg, seed), torch.float32)
    ts.begin_epoch(0)
    n_need = (steps or 64) + warmup
    X, Y = synthetic_ids(cc, min(n_need, 8), data_seed)
    S = c["seq_length"]
    nbm = X.shape[1] // S

    def batch_k(k):
        k %= nbm
        return X[:, k * S:(k + 1) * S], Y[:, k * S:(k + 1) * S]
    losses = []
    for k in range(warmup):
        losses.append(ts.step(*batch_k(k)))
    t0 = time.perf_counter()
    done = 0
    loss = None
    while True:
        loss = ts.step(*batch_k(warmup + done))
        losses.append(loss)
        done += 1
        el = time.perf_counter() - t0
        if steps is not None and done >= steps:
            break
        if steps is None and (el >= max_seconds or done >= 64):
            break
    el = time.perf_counter() - t0
    return dict(value=B * S * done / el, cores=cores, steps=done, seconds=el, batch=B, loss=loss, losses=losses)


def cpu_worker(argv):
    """Hidden mode (``--_cpu_worker``): one CPU replica of the multi-worker reference arm.  Pinned to its own core
    set, it times K steps of the restatement on ITS shard's batches and prints one JSON line."""
    cname, r, n, bs, threads, steps, warmup = argv[0], int(argv[1]), int(argv[2]), int(argv[3]), int(argv[4]), int(argv[5]), int(argv[6])
    ncpu = os.cpu_count() or 1
    per = max(1, ncpu // n)
    try:
        os.sched_setaffinity(0, set(range(r * per, min(ncpu, (r + 1) * per))))
    except Exception:
        pass
    res = cpu_restatement_rate(CONFIGS[cname], 1e9, warmup, steps=steps, batch=bs, threads=threads, seed=0, data_seed=1234 + r)
    print(json.dumps({"rank": r, "seconds": res["seconds"], "steps": res["steps"], "batch": bs, "threads": threads, "loss": res["loss"]}), flush=True)


def run_reference(args, c, rank, world, emit):
    """--impl reference: the reference's CPU path (restated; TF 1.8 is not installable).

    N = 1: one process on the host threads ``cpu_plan`` picks.  N > 1 (``--gpus N``): the reference's distributed
    mode is N worker processes (launch.sh:1-3), so rank 0 starts N CPU workers on disjoint core sets, each stepping
    on its own shard's batches, and reports the aggregate - the same total work per step as the N-GPU arm.  (The
    parameter-server exchange of the reference is left out, which can only flatter the CPU arm.)"""
    if rank != 0:
        return
    n = max(1, args.gpus)
    S = c["seq_length"]
    total = max(1, args.steps + args.warmup)
    plan = cpu_plan(c, per_step_budget=min(6.0, 120.0 / total))
    bs = plan["batch"]
    if n == 1:
        r = cpu_restatement_rate(c, 1e9, args.warmup, steps=args.steps, batch=bs, threads=plan["threads"])
        value, seconds, cores, per_replica = r["value"], r["seconds"], r["cores"], None
    else:
        ncpu = os.cpu_count() or 1
        th = max(1, min(plan["threads"], ncpu // n))
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--_cpu_worker", args.config, str(r), str(n), str(bs), str(th),
                                   str(args.steps), str(args.warmup)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True,
                                  env=dict(os.environ, OMP_NUM_THREADS=str(th), MKL_NUM_THREADS=str(th)))
                 for r in range(n)]
        outs = [json.loads(p.communicate()[0].strip().splitlines()[-1]) for p in procs]
        seconds = max(o["seconds"] for o in outs)
        value = n * bs * S * args.steps / seconds
        cores, per_replica = th * n, [round(bs * S * o["steps"] / o["seconds"], 1) for o in outs]
    sample = "%d warm-up + %d steps of [%d of %d sequences x S=%d] per step%s" % (
        args.warmup, args.steps, bs, c["batch_size"], S, "" if n == 1 else ", %d worker processes on disjoint core sets" % n)
    line = {"impl": "reference", "metric": "train_chars_per_sec", "value": value, "unit": "chars/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * seconds / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": c["desc"], "name": args.config, "host_threads": cores, "host_cpus": os.cpu_count(),
                       "workers": n, "host_threads_probe": plan},
            "cpu_baseline": {"value": value, "unit": "chars/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "chars/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "per_replica": per_replica,
            "note": "TF 1.8 (requirements.txt:2) is not installable here; this is the PyTorch-CPU restatement of "
                    "model.py:8-98 op for op (oracle/torch_restatement.py) on the host cores"}
    emit(line)


def default_precision(cname):
    return "bf16" if cname == "c4" else "bf16x3"


PRECISION_NOTE = {"fp32": "fp32 FFMA contractions, 1e-4 parity",
                  "bf16": "bf16 tcgen05 contractions, fp32 accumulate and state, 1e-2 parity",
                  "bf16x3": "tcgen05 contractions on split-bf16 operands (hi*hi + lo*hi + hi*lo), fp32 accumulate and state: "
                            "fp32-grade, 1e-4 parity"}


def measure_config(cname, precision, K, W, env, want_e2e=True, nprof=10):
    """Device-resident throughput, end-to-end throughput and the live per-phase / per-kernel profile of ONE config.

    ``env``: dict(rank, world, local_rank, pg, lib, torch, dist, Model, _lib).  Returns the pieces of the JSON line."""
    torch, dist, lib, _lib, Model = env["torch"], env["dist"], env["lib"], env["_lib"], env["Model"]
    rank, world, local_rank, pg = env["rank"], env["world"], env["local_rank"], env["pg"]
    c = CONFIGS[cname]
    margs = argparse.Namespace(grad_clip=5.0, output_keep_prob=1.0, input_keep_prob=1.0,
                               **{k: v for k, v in c.items() if k != "desc"})
    model = Model(margs, device="cuda:%d" % local_rank, seed=0, precision=precision, process_group=pg)
    model.assign_lr(0.002)
    B, S = c["batch_size"], c["seq_length"]
    nb = W + K
    X, Y = synthetic_ids(c, nb, 1234 + rank)                  # one shard per rank (data_splitter contract)
    Xd, Yd = torch.from_numpy(X).cuda(), torch.from_numpy(Y).cuda()

    def dev_batch(k):
        return Xd[:, k * S:(k + 1) * S], Yd[:, k * S:(k + 1) * S]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ws_bytes = model.workspace.numel() + 4 * 4 * model.P
    flush = ws_bytes < 2 * L2_BYTES
    flush_buf = torch.empty(int(256e6), dtype=torch.uint8, device="cuda") if flush else None

    # ---- device-resident throughput ------------------------------------------------------
    for k in range(W):
        model.step_async(*dev_batch(k))
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = lib.crnn_launch_count()
    host_s = 0.0
    if flush:
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        for k in range(K):
            flush_buf.zero_()
            evs[k][0].record()
            t0 = time.perf_counter()
            model.step_async(*dev_batch(W + k))
            host_s += time.perf_counter() - t0
            evs[k][1].record()
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
    else:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(K):
            t0 = time.perf_counter()
            model.step_async(*dev_batch(W + k))
            host_s += time.perf_counter() - t0
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
    launches = lib.crnn_launch_count() - l0
    clocks = sampler.summary()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    final_loss = float(model.grads[model.P + 1]) / world
    value = world * B * S * K / (ms * 1e-3)

    # ---- end to end: host ids in, loss out, every step -----------------------------------
    e2e = None
    if want_e2e:
        for k in range(2):
            model.step(X[:, k * S:(k + 1) * S], Y[:, k * S:(k + 1) * S])
        barrier()
        t0 = time.perf_counter()
        for k in range(K):
            model.step(X[:, (W + k) * S:(W + k + 1) * S], Y[:, (W + k) * S:(W + k + 1) * S])
        barrier()
        el = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(el, op=dist.ReduceOp.MAX)
        e2e = {"value": world * B * S * K / float(el.item()), "unit": "chars/s",
               "h2d_bytes_per_step": 2 * B * S * 4, "d2h_bytes_per_step": 4}

    # ---- per-phase timing of the same step (CUDA events on the launching stream) ---------
    pk = peaks()
    roofline = phases = None
    nprof = max(1, min(K, nprof))
    arr = (ctypes.c_double * _lib.NUM_PHASES)()
    if rank == 0:
        _lib.check(lib.crnn_profile_enable(model._h, 1))
    kms = (ctypes.c_double * 2)()
    kln = (ctypes.c_int64 * 2)()
    for k in range(nprof):                                    # every rank steps (the all-reduce is collective)
        model.step_async(*dev_batch(W + k))
        if rank == 0:
            _lib.check(lib.crnn_profile_read(model._h, arr, 0))
            _lib.check(lib.crnn_profile_read_kernels(model._h, kms, kln, 0))
    barrier()
    if rank == 0:
        _lib.check(lib.crnn_profile_enable(model._h, 0))
        phases = {lib.crnn_profile_phase_name(i).decode(): arr[i] / nprof for i in range(_lib.NUM_PHASES)}
        roofline = make_roofline(c, cname, precision, phases, kms, kln, nprof, model.P, pk)
        step_tf = value / world * train_flop_per_char(c) / 1e12
        roofline["step_tflops"] = step_tf
        roofline["step_frac_of_tensor_peak"] = step_tf / pk["tf_sus"]
        roofline["hbm_phases"] = hbm_phases(c, phases, model.P, pk, 2 if precision == "bf16x3" else 1)
    out = dict(value=value, ms_per_step=ms / K, launches=int(launches), clocks=clocks, e2e=e2e, roofline=roofline,
               phases_ms=phases, final_loss=final_loss, host_ms_per_step=1e3 * host_s / K,
               l2=("256 MB flush write between timed steps (excluded from timing)" if flush else
                   "per-step working set %.0f MB > 126 MB L2, new batch every step" % (ws_bytes / 1e6)),
               margs=margs, dev_batch=dev_batch, nb=nb)
    del model
    torch.cuda.empty_cache()
    return out


def hbm_phases(c, phases, P, pk, nsp=2):
    """Achieved GB/s of the bandwidth-bound phases (algorithmic bytes / live phase time) against the measured copy peak."""
    by = phase_bytes(c, P)
    if phases.get("softmax_xent", 0.0) < 0.005 and phases.get("logits_gemm", 0.0) > 0:
        # fused projection + softmax + cross-entropy (csrc/proj_xent_tc.cuh): the logits never reach HBM; the phase reads the
        # top layer's bf16 rows (hi + lo) and writes bf16 dlogits (hi + lo)
        SB, Vp = c["seq_length"] * c["batch_size"], (c["vocab_size"] + 7) // 8 * 8
        by.pop("softmax_xent", None)
        by["logits_gemm"] = SB * c["rnn_size"] * 2 * nsp + SB * Vp * 2 * nsp + SB * 4
    out = {}
    if phases.get("xproj_gemm", 0.0) < 1e-6 and c["num_layers"] > 1 and "embed_scatter" in by:
        # layer wavefront: the sorted embedding scatter runs on a side stream beside the last weight-gradient GEMMs (crnn_api.cu:
        # early_scatter); what is left of its phase on the main stream is one scalar kernel, not the bytes - no live figure
        # (profiles/r02_ncu_hbm_c3.txt has the kernel alone: 52.7 MB in 23.4 us cold)
        b = by.pop("embed_scatter")
        out["embed_scatter"] = {"ms": None, "algorithmic_bytes": b, "achieved_gbs": None, "frac": None,
                                "note": "overlapped with the weight-gradient GEMMs on a side stream; see profiles/r02_ncu_hbm_c3.txt"}
    for name, b in by.items():
        ms = phases.get(name, 0.0)
        if ms > 1e-6:
            gbs = b / (ms * 1e-3) / 1e9
            out[name] = {"ms": round(ms, 4), "algorithmic_bytes": b, "achieved_gbs": round(gbs, 1), "frac": round(gbs / pk["hbm"], 4)}
    return out


def make_roofline(c, cname, precision, phases, kms, kln, nprof, P, pk):
    fl, by = phase_flops(c), phase_bytes(c, P)
    if phases.get("xproj_gemm", 0.0) < 1e-6 and c["num_layers"] > 1:
        # layer wavefront: the recurrence phases are fork..join spans that contain the interleaved
        # x-projection / input- and weight-gradient GEMMs, so they carry those FLOPs too
        fl["rec_fwd"] += fl["xproj_gemm"]
        fl["rec_bwd"] += fl["input_grads"] + fl["weight_grads"]
    dom = max((p for p in phases if p != "misc"), key=lambda p: phases[p])
    dms = phases[dom]
    if dom in ("rec_fwd", "rec_bwd") and kln[1 if dom == "rec_bwd" else 0] > 0:
        # the persistent recurrence kernel itself: algorithmic FLOPs of all its launches in one step
        # over the summed durations of those launches (events on each launch's own stream)
        i = 1 if dom == "rec_bwd" else 0
        per_step_ms = kms[i] / nprof
        kflops = phase_flops(c)[dom]
        ach = kflops / (per_step_ms * 1e-3) / 1e12
        roofline = {"bound": "tensor", "kernel": "rec_%s_tc_kernel" % ("bwd" if i else "fwd"), "achieved": ach,
                    "peak": pk["tf_sus"], "unit": "TFLOP/s", "frac": ach / pk["tf_sus"], "traffic": None,
                    "peak_source": pk["src"] + " bf16 sustained", "launches_per_step": kln[i] / nprof,
                    "ms": kms[i] / max(1, kln[i]), "ms_per_step_summed": per_step_ms,
                    "algorithmic_flops_per_launch": kflops / (kln[i] / nprof),
                    "tensor_products_per_algorithmic_product": 3 if precision == "bf16x3" else 1,
                    "phase_span_ms": dms}
    elif dom in fl:
        ach = fl[dom] / (dms * 1e-3) / 1e12
        roofline = {"bound": "tensor", "kernel": dom, "achieved": ach, "peak": pk["tf_sus"], "unit": "TFLOP/s",
                    "frac": ach / pk["tf_sus"], "traffic": None, "peak_source": pk["src"] + " bf16 sustained",
                    "ms": dms, "algorithmic_flops": fl[dom]}
    else:
        ach = by[dom] / (dms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": pk["hbm"], "unit": "GB/s",
                    "frac": ach / pk["hbm"], "traffic": None, "peak_source": pk["src"], "ms": dms,
                    "algorithmic_bytes": by[dom]}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):      # dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed ncu capture
        roofline["traffic"] = json.load(open(tpath)).get(cname, {}).get(dom)
    return roofline


def loss_parity(cname, precision, env, batch, steps, threads):
    """The same first steps (same seed-0 weights, same batches) on the CPU restatement and on the GPU path, at ``batch``
    sequences per step (the full batch where one CPU step is affordable, a B/8 slice on both arms for configs 4/5)."""
    torch, Model = env["torch"], env["Model"]
    c = dict(CONFIGS[cname], batch_size=batch)
    S = c["seq_length"]
    r = cpu_restatement_rate(CONFIGS[cname], 1e9, 0, steps=steps, batch=batch, threads=threads)
    margs = argparse.Namespace(grad_clip=5.0, output_keep_prob=1.0, input_keep_prob=1.0, **{k: v for k, v in c.items() if k != "desc"})
    m2 = Model(margs, device="cuda:%d" % env["local_rank"], seed=0, precision=precision)
    m2.assign_lr(0.002)
    X0, Y0 = synthetic_ids(c, min(steps, 8), 1234)
    nbm = X0.shape[1] // S
    gl = [m2.step(X0[:, (k % nbm) * S:(k % nbm + 1) * S], Y0[:, (k % nbm) * S:(k % nbm + 1) * S])[0] for k in range(steps)]
    rel = max(abs(a_ - b_) / abs(b_) for a_, b_ in zip(gl, r["losses"]))
    del m2
    torch.cuda.empty_cache()
    return {"steps": steps, "batch": batch, "max_rel_loss_diff": rel, "gpu": gl, "cpu": r["losses"][:steps],
            "tolerance": 2e-2 if precision == "bf16" else 1e-4}, r


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--_cpu_worker":
        cpu_worker(sys.argv[2:])
        return
    # Only the JSON line may reach stdout: libraries (NCCL's version banner, torchrun notices)
    # print there too, so stdout is pointed at stderr until the line is ready.
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)

    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default=os.environ.get("CRNN_BENCH_CONFIG", "c3"), choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=None, choices=["fp32", "bf16", "bf16x3"],
                    help="default: bf16x3 (fp32-grade split-bf16 tensor-core contractions, 1e-4 parity) except config 4, "
                         "which BASELINE.json specifies as bf16 gate GEMMs")
    ap.add_argument("--no-alt", action="store_true", help="skip the short runs of the other precisions")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the short runs of the other BASELINE.json configs")
    args = ap.parse_args()
    c = CONFIGS[args.config]
    if args.precision is None:
        args.precision = default_precision(args.config)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, c, rank, world, emit)
        return
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (no CPU fallback); use --impl reference for the CPU arm")
    ge.build()
    from distributed_char_rnn_b200 import _lib
    from distributed_char_rnn_b200.model import Model
    torch.cuda.set_device(local_rank)
    pg = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        pg = dist.group.WORLD
    lib = _lib.load()
    env = dict(rank=rank, world=world, local_rank=local_rank, pg=pg, lib=lib, torch=torch, dist=dist, Model=Model, _lib=_lib)
    B, S = c["batch_size"], c["seq_length"]
    K, W = args.steps, args.warmup

    main_run = measure_config(args.config, args.precision, K, W, env, want_e2e=not args.no_e2e)
    margs, dev_batch, nb = main_run["margs"], main_run["dev_batch"], main_run["nb"]

    # ---- the other precisions of the same step, short device-resident runs (context for the headline) ----
    alt = None
    if rank == 0 and world == 1 and not args.no_alt:
        alt = {}
        for prec in ("fp32", "bf16", "bf16x3"):
            if prec == args.precision:
                continue
            try:
                m3 = Model(margs, device="cuda:%d" % local_rank, seed=0, precision=prec)
            except _lib.CrnnError as ex:
                alt[prec] = {"unsupported": str(ex)[:120]}
                continue
            m3.assign_lr(0.002)
            n3 = 3 if prec == "fp32" else min(K, 10)
            for k in range(3):
                m3.step_async(*dev_batch(k))
            torch.cuda.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            for k in range(n3):
                m3.step_async(*dev_batch((W + k) % nb))
            a1.record()
            torch.cuda.synchronize()
            t3 = a0.elapsed_time(a1) / n3
            alt[prec] = {"ms_per_step": t3, "value": B * S / (t3 * 1e-3), "unit": "chars/s", "steps": n3,
                         "parity_tolerance": 2e-2 if prec == "bf16" else 1e-4}
            del m3
            torch.cuda.empty_cache()

    if world > 1:
        dist.barrier()
    cpu = parity = plan = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # bound the sample: a CPU step of the large configs is tens of TFLOP; use fewer sequences per step
        plan = cpu_plan(c, per_step_budget=3.0)
        bs = plan["batch"]
        parity, r0 = loss_parity(args.config, args.precision, env, bs, 4 if bs == B else 2, plan["threads"])
        r = cpu_restatement_rate(c, 12.0, 1, batch=bs, threads=plan["threads"])
        cpu = {"value": r["value"], "unit": "chars/s", "cores": r["cores"], "kind": "port", "host_threads_probe": plan,
               "sample": "1 warm-up + %d steps of [%d of %d sequences x S=%d] (%d chars each) in %.1f s"
                         % (r["steps"], bs, B, S, bs * S, r["seconds"])}

    # ---- the other configs of BASELINE.json: short runs, each with its own clocks, roofline and loss parity ----
    others = None
    if rank == 0 and world == 1 and not args.no_other_configs:
        others = {}
        short = {"c1": 50, "c3": 20, "c4": 10, "c5": 5}
        for oc in sorted(CONFIGS):
            if oc == args.config:
                continue
            try:
                prec = default_precision(oc)
                mr = measure_config(oc, prec, short[oc], 3, env, want_e2e=not args.no_e2e, nprof=3)
                entry = {"workload": CONFIGS[oc]["desc"], "precision": prec, "value": mr["value"], "unit": "chars/s",
                         "ms_per_step": mr["ms_per_step"], "steps": short[oc], "warmup": 3, "e2e": mr["e2e"], "clocks": mr["clocks"],
                         "roofline": mr["roofline"], "gpu_launches": mr["launches"], "host_ms_per_step": mr["host_ms_per_step"],
                         "phases_ms": mr["phases_ms"], "l2": mr["l2"], "final_loss": mr["final_loss"]}
                if not args.no_cpu_baseline:
                    ob = CONFIGS[oc]["batch_size"]
                    pb = ob if oc in ("c1", "c3") else max(1, ob // 8)
                    entry["loss_parity"], rr = loss_parity(oc, prec, env, pb, 4 if pb == ob else 2, plan["threads"] if plan else None)
                    entry["cpu_baseline"] = {"value": rr["value"], "unit": "chars/s", "cores": rr["cores"], "kind": "port",
                                             "sample": "%d steps of [%d of %d sequences x S=%d]" % (rr["steps"], pb, ob, CONFIGS[oc]["seq_length"])}
                if oc == "c4" and not args.no_alt:
                    # config 4 is quoted in bf16 (BASELINE.json's own wording); the same step at the fp32 parity bar
                    x3 = measure_config(oc, "bf16x3", 5, 3, env, want_e2e=False, nprof=0)
                    entry["other_precisions"] = {"bf16x3": {"ms_per_step": x3["ms_per_step"], "value": x3["value"], "unit": "chars/s",
                                                            "steps": 5, "parity_tolerance": 1e-4}}
                others[oc] = entry
            except Exception as ex:                       # a failing side run must not cost the headline line
                others[oc] = {"error": str(ex)[:200]}

    if rank == 0:
        line = {"metric": "train_chars_per_sec", "value": main_run["value"], "unit": "chars/s", "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": main_run["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": {"fp32": "f32", "bf16": "bf16", "bf16x3": "bf16x3"}[args.precision], "data": "synthetic",
                "config": {"workload": c["desc"], "name": args.config, "global_batch": world * B, "seq_length": S,
                           "parallelism": "dp%d" % world, "precision": args.precision,
                           "precision_note": PRECISION_NOTE[args.precision], "l2": main_run["l2"]},
                "clocks": main_run["clocks"], "e2e": main_run["e2e"], "gpu_launches": main_run["launches"],
                "host_ms_per_step": main_run["host_ms_per_step"], "roofline": main_run["roofline"],
                "cpu_baseline": cpu, "loss_parity": parity, "phases_ms": main_run["phases_ms"], "final_loss": main_run["final_loss"],
                "other_precisions": alt, "other_configs": others}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
