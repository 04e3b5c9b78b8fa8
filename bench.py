#!/usr/bin/env python
"""
bench.py — headline benchmark of the Grappa hot path on B200 (contract: see the task statement / DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (one process per GPU under torchrun for N > 1)
  python bench.py --impl reference [--steps K] [--warmup W]       the reference's algorithm on the host cores (oracle port)

Workload at N = 1: BASELINE.json configs[1] — one training step (GrappaModel -> Energy -> molecule-wise loss ->
backward -> AdamW, dropout active) on a synthetic SPICE-like batch of 32 molecules x 40 atoms x 50 conformations.
For N > 1 every rank owns its own 32-molecule slice (weak scaling) and gradients are all-reduced over NCCL.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

# stdout must carry exactly ONE JSON line: park the real stdout and point fd 1 at stderr, so that anything a library
# prints there (e.g. NCCL's version banner) cannot get in front of the result
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line: dict):
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


import torch  # noqa: E402

MOLS_PER_RANK = 32
N_CONFS = 50
WORKLOAD = ("cfg2: training step (fwd + molecule-wise energy/force loss + bwd + AdamW, dropout on), "
            "32 molecules x 40 atoms x 50 conformations per GPU, default Grappa model (20.8 M params)")


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return d, "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def __enter__(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0=None, t1=None):
        """Samples taken inside [t0, t1] (the timed regions; nvidia-smi is started before the warm-up because its start-up
        alone outlasts a 0.1 s timed region)."""
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        lines = [ln for ts_, ln in self.lines if (t0 is None or ts_ >= t0) and (t1 is None or ts_ <= t1 + 0.12)]
        for ln in lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def build_batch(first_seed: int, n_mol: int = MOLS_PER_RANK, n_confs: int = N_CONFS):
    from grappa_b200 import synthetic as S
    return S.cfg2_batch(n_mol, n_confs, first_seed)


def cpu_train_baseline(sd, g, steps: int, warmup: int, budget_s: float = 25.0):
    """Oracle training step (dropout on, AdamW) on all host threads; bounded sample of the same workload: the step
    count is honoured unless steps x (time of one step) would exceed `budget_s`, then fewer steps are timed (said in
    `sample`); below 8 molecules per step the sample would stop being the same workload, so that is the floor."""
    import grappa_oracle as O
    from grappa_b200.graph import batch, unbatch
    torch.set_num_threads(os.cpu_count() or 1)
    torch.set_float32_matmul_precision("highest")
    tr = O.OracleTrainer(sd)
    n_mol = g.batch_size
    t0 = time.perf_counter()
    tr.step(g)
    t_first = time.perf_counter() - t0
    sample_mols = n_mol
    if t_first * (min(steps, 3) + max(warmup - 1, 0)) > budget_s and n_mol > 8:
        sample_mols = 8
        g = batch(unbatch(g)[:8])          # unbatch() returns views with local idxs; re-batch the first 8 molecules
    for _ in range(max(warmup - 1, 0)):
        tr.step(g)
    t0 = time.perf_counter()
    done = 0
    while done < steps:
        tr.step(g)
        done += 1
        if time.perf_counter() - t0 > budget_s and done >= 3:
            break
    dt = (time.perf_counter() - t0) / done
    return {"value": sample_mols / dt, "unit": "molecules/s", "cores": torch.get_num_threads(), "kind": "port",
            "steps_run": done,
            "sample": f"{done} training steps (+{warmup} warm-up) of {sample_mols} molecules x 40 atoms x {N_CONFS} conformations, "
                      f"oracle port of the reference (PyTorch CPU fp32, dropout on, AdamW), {dt * 1e3:.0f} ms/step"}, dt


def _reference_state_dict():
    """Random-init weights with the reference's 296-key layout, WITHOUT importing the product's model code (and with
    it libgrappa_b200.so) into the CPU arm: shapes from tests/golden/state_dict_layout.json (written from the
    reference's own GrappaModel by oracle/make_golden.py), its non-learnable buffers from state_dict_constants.npz (same
    script), learnable values from the oracle's deterministic fill."""
    from weights import deterministic_fill_
    with open(os.path.join(ROOT, "tests", "golden", "state_dict_layout.json")) as f:
        layout = json.load(f)
    import numpy as np
    sd = {k: torch.zeros(tuple(shape), dtype=getattr(torch, dtype.replace("torch.", ""))) for k, (shape, dtype) in layout.items()}
    consts = np.load(os.path.join(ROOT, "tests", "golden", "state_dict_constants.npz"))   # statistics, permutations, encodings
    for k in consts.files:
        sd[k] = torch.as_tensor(consts[k]).to(sd[k].dtype).reshape(sd[k].shape)
    deterministic_fill_(sd)
    return sd


def run_reference(args):
    """--impl reference: the reference's algorithm (oracle port; /root/reference cannot travel and needs DGL) on all host
    cores.  Rank 0 only.  The metric is a RATE (molecules/s): every step is one 32-molecule cfg-2 batch whatever --gpus
    says (a CPU process gains nothing from a larger batch), so the rate is comparable with the product's aggregate."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sd = _reference_state_dict()
    g = build_batch(0)
    steps, warmup = max(1, args.steps), max(1, args.warmup)
    base, dt = cpu_train_baseline(sd, g, steps=steps, warmup=warmup, budget_s=240.0)
    line = {"impl": "reference", "metric": "train_molecules_per_s", "value": base["value"], "unit": "molecules/s",
            "n_gpus": args.gpus, "steps": base["steps_run"], "warmup": warmup, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "global_batch": MOLS_PER_RANK,
                       "note": "CPU oracle port on rank 0; a rate over 32-molecule steps, independent of --gpus"},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "molecules/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the inference / energy-sweep side measurements")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="eager launches instead of CUDA-graph replay")
    ap.add_argument("--no-split-update", action="store_true",
                    help="one all-reduce + one AdamW after backward instead of the GNN / writers split (A/B switch)")
    ap.add_argument("--split-update", action="store_true", help="staged update under the next forward also on one GPU (A/B switch)")
    ap.add_argument("--math-mode", type=int, default=None)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    from grappa_b200 import _lib, ops
    from grappa_b200.batching import PackedBatch
    from grappa_b200.model import GrappaModel
    from grappa_b200.train_step import TrainStep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)
    if args.math_mode is not None:
        ops.MATH_MODE = args.math_mode
    K, W = args.steps, max(args.warmup, 3)

    torch.manual_seed(0)                       # identical random-init weights on every rank
    model = GrappaModel.from_default_config()
    sd_cpu = {k: v.clone() for k, v in model.state_dict().items()}
    ts = TrainStep(model, dev, lr=1e-5, weight_decay=0.0, world_size=world, use_cuda_graph=not args.no_graph,
                   split_update=False if args.no_split_update else (True if args.split_update else None))
    ops.DropoutSeeds.manual_seed(1234 + rank)  # dropout streams differ per rank
    g_host = build_batch(first_seed=MOLS_PER_RANK * rank)
    packed = PackedBatch(g_host, pin=True)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local)
    clocks.__enter__()
    # ---- warm-up (includes CUDA-graph capture after two eager steps) ----
    l0 = _lib.launch_count()
    ts.step(packed, load=True)
    torch.cuda.synchronize()
    launches_per_step = _lib.launch_count() - l0
    for _ in range(max(W - 1, 3)):
        ts.step(packed, load=True)
    barrier()

    # ---- set-up of the end-to-end region (molecule store, its batch shape captured), before any timing ----
    import numpy as np
    from grappa_b200 import synthetic as S
    from grappa_b200.store import MoleculeStore
    store = MoleculeStore(S.cfg2_graphs(3 * MOLS_PER_RANK, N_CONFS, first_seed=1000 + 3 * MOLS_PER_RANK * rank), dev)
    rng = np.random.default_rng(rank)
    draws = [rng.choice(store.n, MOLS_PER_RANK, replace=False) for _ in range(K + 6)]
    for ids in draws[:6]:                              # warm-up incl. CUDA-graph capture for this batch shape
        ts.step_from_store(store, ids)
    h2d_e2e = store.prepare(draws[0]).h2d_bytes
    # ---- timed region 1: device-resident inputs ----
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_clock0 = time.time()
    start.record()
    for _ in range(K):
        loss = ts.step(packed, load=False)
    end.record()
    barrier()
    ms = start.elapsed_time(end)
    # ---- timed region 2: END TO END through the public training API with a DIFFERENT batch every step ----
    # Every step draws 32 molecules from a device-resident molecule store (store.MoleculeStore):
    # the host work of a real loop (store.prepare: conformation choice, segment tables), the host->device copy of that
    # step's inputs (the pinned segment tables; the molecules' own data already live in HBM) and the device->host read
    # of the step's loss are all inside the timed region.  The loss of step k is consumed while step k+1 is being
    # issued, as a training loop that logs its loss does.
    barrier()
    start2, end2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start2.record()
    last, pending = 0.0, None
    for ids in draws[6:]:
        loss = ts.step_from_store(store, ids)
        handle = ts.read_loss_async(loss)
        if pending is not None:
            last = pending.value()
        pending = handle
    last = pending.value()
    end2.record()
    barrier()
    ms2 = start2.elapsed_time(end2)
    # ---- the same batch replayed from pinned host memory (2.8 MB H2D per step, prefetched on a copy stream): what the
    #      round-1 `e2e` measured; kept as a side figure ----
    barrier()
    start3, end3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts.prefetch(packed)
    start3.record()
    pending = None
    for _ in range(K):
        loss = ts.step(packed, load=True)
        ts.prefetch(packed)
        handle = ts.read_loss_async(loss)
        if pending is not None:
            pending.value()
        pending = handle
    pending.value()
    end3.record()
    barrier()
    ms3 = start3.elapsed_time(end3)
    t_clock1 = time.time()
    clocks.__exit__()
    t = torch.tensor([ms, ms2, ms3], device=dev, dtype=torch.float64)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    ms, ms2, ms3 = float(t[0]), float(t[1]), float(t[2])
    mols = MOLS_PER_RANK * world

    line = {"metric": "train_molecules_per_s", "value": mols * K / (ms * 1e-3), "unit": "molecules/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "global_batch": mols, "parallelism": f"dp{world}",
                       "gemm_math": {0: "fp32 FFMA", 1: "tcgen05 3xTF32", 2: "tcgen05 bf16x3 (forward, dX) + 3xTF32 (dW)"}[ops.MATH_MODE],
                       "cuda_graph": not args.no_graph,
                       "l2": "no explicit flush: one step touches parameters + gradients + Adam state (4 x 83 MB) plus "
                             "~0.5 GB of activations, several times the 126 MB L2"},
            "e2e": {"value": mols * K / (ms2 * 1e-3), "unit": "molecules/s", "h2d_bytes_per_step": h2d_e2e,
                    "d2h_bytes_per_step": 4, "ms_per_step": ms2 / K,
                    "what": "TrainStep.step_from_store: a different 32-molecule batch per rank every step, host batch "
                            "preparation + H2D of its tables + device collate + step + loss read-back inside the timed region",
                    "replayed_packed_batch": {"value": mols * K / (ms3 * 1e-3), "ms_per_step": ms3 / K,
                                              "h2d_bytes_per_step": packed.h2d_bytes}},
            "gpu_launches": int(launches_per_step * K), "final_loss": last}
    if rank == 0:
        line["clocks"] = clocks.summary(t_clock0, t_clock1)
        line["clocks"]["window"] = "device-resident + both end-to-end timed regions (back to back, all under load)"
    if world > 1:
        line["config"]["gradient_exchange"] = ("grappa_dp_exchange_adamw: in-switch gradient sum (multimem.ld_reduce) + AdamW on the "
                                               "owner's shard + multicast store of the parameters, one kernel per stage, stages "
                                               "after the first under the next forward" if ts._switch is not None else
                                               "NCCL all-reduce per stage + AdamW on every rank, stages after the first under the next forward")
        line["dp_check"] = dp_check(ts, world, rank, dev)
    line["strong_scaling_global256"] = strong_scaling(ts, world, rank, dev, K)

    # ---- per-kernel timing with CUDA events (eager pass, rank 0) -> roofline ----
    if rank == 0:
        pk, pk_src = peaks()
        ts_graph, ts_switch = ts.use_cuda_graph, ts._switch
        try:
            _lib.PROFILE_HINTS.update(n_edges=g_host.num_edges(), n_atoms=g_host.num_nodes("n1"))
            ts.use_cuda_graph = False
            # this pass runs on rank 0 ALONE: the update must not wait for the other ranks, so it is the local AdamW kernel
            # (at N > 1 the timed steps use grappa_dp_exchange_adamw instead, which needs every rank)
            ts._switch = None
            for _ in range(2):
                ts.compute_gradients(packed, load=False)
            prof = _lib.Profiler()
            _lib.profiler = prof
            ts.compute_gradients(packed, load=False)      # ONE step's worth of C-ABI calls, with their arguments
            ts.optimizer_step()
            summ = prof.summary()                          # call counts and algorithmic work per entry point
            _lib.profiler = None
            ts.use_cuda_graph = ts_graph
            ts._switch = ts_switch
            # device time per entry point: its recorded launches replayed back to back in one CUDA graph (CUDA events)
            gemm_names = {"grappa_gemm", "grappa_gemm_actgrad", "grappa_gemm_ex", "grappa_gemm_bf16x3"}     # one kernel family: one roofline entry
            summ.setdefault("grappa_gemm", {"calls": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0})
            for other in ("grappa_gemm_actgrad", "grappa_gemm_ex", "grappa_gemm_bf16x3"):
                if other in summ:
                    extra_g = summ.pop(other)
                    for k2 in ("calls", "ms", "flops"):
                        summ["grappa_gemm"][k2] += extra_g[k2]
            names = sorted(summ.keys())
            dev_ms = {n: prof.replay_ms(gemm_names if n == "grappa_gemm" else {n}) for n in names}
            tot = sum(dev_ms.values())
            step_ms = ms / K
            gm = summ["grappa_gemm"]
            gemm_ms = dev_ms["grappa_gemm"]
            tf = gm["flops"] / (gemm_ms * 1e-3) / 1e12
            peak_tf = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
            # fp32-accurate ceiling: a split product costs 3 MMAs; kind::f16 (bf16 planes) runs at the bf16 peak, kind::tf32 at
            # half of it -> time floor = flops_bf16x3 / (peak / 3) + flops_3xtf32 / (peak / 6)
            _mnk = lambda n, a: (a[0], a[1], a[2]) if n == "grappa_gemm_bf16x3" else (a[1], a[2], a[3])
            _gflops = lambda keep: (sum(2.0 * _mnk(n, a)[0] * _mnk(n, a)[1] * _mnk(n, a)[2] for n, _, _, a in prof.records
                                        if n == "grappa_gemm_bf16x3" and keep(n, a)),
                                    sum(2.0 * _mnk(n, a)[0] * _mnk(n, a)[1] * _mnk(n, a)[2] for n, _, _, a in prof.records
                                        if n in gemm_names and n != "grappa_gemm_bf16x3" and keep(n, a)))
            def ceiling_frac(keep, t_ms):
                fb, ft = _gflops(keep)
                return (fb / (peak_tf / 3.0) + ft / (peak_tf / 6.0)) / 1e12 / (t_ms * 1e-3)
            every = lambda n, a: True
            line["roofline"] = {"kernel": {0: "sgemm_kernel", 1: "gemm_tc_kernel", 2: "gemm_bf16_kernel + gemm_tc_kernel"}[ops.MATH_MODE] +
                                          " (grappa_gemm: all dense layers fwd+bwd, fused bias/ELU/dropout/residual epilogue)",
                                "bound": "tensor", "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": tf / peak_tf,
                                "frac_of_fp32_accurate_ceiling": ceiling_frac(every, gemm_ms),
                                "flops_bf16x3": _gflops(every)[0], "flops_3xtf32": _gflops(every)[1],
                                "traffic": None, "peak_source": f"{pk_src} bf16 sustained (MEASURED_PEAKS.json)",
                                "pipe": "fp32 FFMA (math_mode 0)" if ops.MATH_MODE == 0 else
                                        "tcgen05, 3 MMAs per product for fp32-level accuracy: kind::f16 on bf16 hi/lo planes (forward, "
                                        "dX; ceiling = bf16 peak / 3) and kind::tf32 (weight gradients; ceiling = bf16 peak / 6); "
                                        "frac_of_fp32_accurate_ceiling = that time floor / measured time",
                                "share_of_step": gemm_ms / tot, "launches_per_step": gm["calls"],
                                "share_note": "share of the summed device time of this library's kernels; the step itself overlaps "
                                              "kernels on parallel streams, so the sum exceeds the step time",
                                "flops_per_step": gm["flops"], "ms_per_step": gemm_ms,
                                "avg_launch_us": gemm_ms * 1e3 / gm["calls"],
                                "method": "the step's recorded launches replayed back to back in a CUDA graph, CUDA events",
                                "in_step": {"achieved": gm["flops"] / (step_ms * 1e-3) / 1e12,
                                            "frac": gm["flops"] / (step_ms * 1e-3) / 1e12 / peak_tf,
                                            "frac_of_fp32_accurate_ceiling": ceiling_frac(every, step_ms),
                                            "note": "the same flops over the measured step time: inside the step the GEMMs of "
                                                    "five streams overlap each other and the other kernels, so this aggregate "
                                                    "rate is what the step sustains; `achieved` above is the serial replay"}}
            # The weight-gradient GEMMs are issued on low-priority side streams with a CTA budget (un-split: 4-64 CTAs
            # each) so that they run UNDER the activation-gradient chains; replayed alone and serially they look slow by
            # construction.  Both classes are reported; `achieved` above is over all launches, as the contract defines it.
            is_bg = lambda n, a: n in ("grappa_gemm", "grappa_gemm_ex", "grappa_gemm_bf16x3") and ((a[19] >> 16) & 1) == 1
            mnk = lambda n, a: (a[0], a[1], a[2]) if n == "grappa_gemm_bf16x3" else (a[1], a[2], a[3])
            fl = {"foreground": [0.0, 0], "background": [0.0, 0]}
            for n, _, _, a in prof.records:
                if n in gemm_names:
                    c = fl["background" if is_bg(n, a) else "foreground"]
                    c[0] += 2.0 * mnk(n, a)[0] * mnk(n, a)[1] * mnk(n, a)[2]
                    c[1] += 1
            for cls, keep in (("foreground", lambda n, a: not is_bg(n, a)), ("background", is_bg)):
                ms_c = prof.replay_ms(gemm_names, keep=keep)
                if ms_c > 0:
                    line["roofline"][cls] = {"launches": fl[cls][1], "flops": fl[cls][0], "ms": ms_c,
                                             "achieved": fl[cls][0] / (ms_c * 1e-3) / 1e12,
                                             "frac_of_fp32_accurate_ceiling": ceiling_frac(keep, ms_c)}
            line["roofline"]["class_note"] = ("foreground = GEMMs on the critical chains (forward, activation gradients); background = "
                                              "weight-gradient GEMMs on low-priority side streams, deliberately un-split (few CTAs) so "
                                              "that they fill idle SMs under the chains: 4.06 -> 3.93 ms per step")
            others = {}
            for name, label in (("grappa_energy_fwd", "energy_fwd"), ("grappa_energy_vjp", "energy_vjp"), ("grappa_gnn_attn_fwd", "gnn_attn_fwd")):
                if name in summ and dev_ms.get(name, 0) > 0:
                    d = summ[name]
                    gbs = d["bytes"] / (dev_ms[name] * 1e-3) / 1e9
                    others[label] = {"bound": "hbm", "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"],
                                     "us_per_call": dev_ms[name] * 1e3 / d["calls"]}
            try:     # DRAM traffic of the dominant kernel from the committed ncu --set full capture (profiles/)
                with open(os.path.join(ROOT, "profiles", "r02_gemm_traffic.json")) as f:
                    tr = json.load(f)
                line["roofline"]["traffic"] = tr["dram_bytes_per_launch"]
                line["roofline"]["traffic_note"] = tr["note"]
            except Exception:
                pass
            line["roofline_other_kernels"] = others
            line["kernel_ms_per_step"] = {k: round(v, 4) for k, v in sorted(dev_ms.items(), key=lambda kv: -kv[1])}
            line["kernel_ms_per_step"]["_sum_own_kernels"] = round(tot, 4)
            line["kernel_ms_per_step"]["_step"] = round(step_ms, 4)
        except Exception as e:        # a failed side measurement must not cost the headline numbers above
            line.setdefault("roofline", {"error": f"kernel profile pass failed: {type(e).__name__}: {str(e)[:300]}"})
            _lib.profiler = None
            ts.use_cuda_graph, ts._switch = ts_graph, ts_switch

    # ---- side measurements on rank 0: inference (cfg 3) and energy sweep (cfg 4) ----
    if rank == 0 and world == 1 and not args.no_extras:
        try:
            line["extra"] = extras(dev, pk)
            line["extra"]["device_collate"] = collate_bench(ts, dev, K)
        except Exception as e:
            line.setdefault("extra", {})["error"] = f"{type(e).__name__}: {str(e)[:300]}"
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            base, _ = cpu_train_baseline(sd_cpu, g_host, steps=3, warmup=1)
            line["cpu_baseline"] = base
        except Exception as e:
            line["cpu_baseline"] = {"error": f"{type(e).__name__}: {str(e)[:300]}"}
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        emit(line)                       # before the last barrier: the other ranks wait there while rank 0 measures
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


def dp_check(ts, world, rank, dev):
    """Correctness of the multi-GPU path, emitted with the bench line (world > 1):
      replicas  after the timed steps every rank must hold bit-identical parameters and Adam second moments (each rank saw
                different molecules and different dropout masks, so only a correct exchange keeps them equal);
      gradient  the all-reduced mean gradient of `world` ranks x 32 molecules equals the gradient one process computes on
                the concatenated 32 x world batch (dropout off, same weights)."""
    from grappa_b200.batching import PackedBatch
    from grappa_b200.model import GrappaModel
    from grappa_b200.train_step import TrainStep
    dist = torch.distributed
    ts.synchronize()
    ts.gather_optimizer_state()            # the fused NVSwitch exchange shards the Adam moments over the ranks
    out = {"exchange": "nvswitch multimem, fused with AdamW (grappa_dp_exchange_adamw)" if ts._switch is not None else "nccl all-reduce"}
    for name, t in (("params", ts.flat.params), ("exp_avg_sq", ts.flat.exp_avg_sq)):
        bits = t.view(torch.int32).to(torch.int64)
        mine = torch.stack([bits.sum(), (bits * (torch.arange(bits.numel(), device=dev) % 8191 + 1)).sum()])
        allv = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allv, mine)
        out[name + "_identical"] = bool(all(torch.equal(v, allv[0]) for v in allv))
    torch.manual_seed(0)
    model = GrappaModel.from_default_config()
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    t2 = TrainStep(model, dev, world_size=world, use_cuda_graph=False, switch_exchange=False)
    loss_local = t2.compute_gradients(PackedBatch(build_batch(MOLS_PER_RANK * rank), pin=False))
    g_dp = t2.flat.grads.clone()
    dist.all_reduce(g_dp)
    g_dp /= world
    l_dp = loss_local.clone().reshape(1)
    dist.all_reduce(l_dp)
    l_dp /= world
    loss_glob = t2.compute_gradients(PackedBatch(build_batch(0, n_mol=MOLS_PER_RANK * world), pin=False))
    g_1 = t2.flat.grads
    out["grad_rel_err"] = float((g_dp - g_1).abs().max() / g_1.abs().max())
    out["loss_rel_err"] = float((l_dp[0] - loss_glob).abs() / loss_glob.abs())
    out["ok"] = bool(out["params_identical"] and out["exp_avg_sq_identical"] and out["grad_rel_err"] < 1e-3 and out["loss_rel_err"] < 1e-4)
    del t2
    return out


def strong_scaling(ts, world, rank, dev, K):
    """BASELINE.json configs[4] as stated: a fixed 256-molecule GLOBAL batch split over the ranks (256 / world each)."""
    from grappa_b200.batching import PackedBatch
    per = 256 // world
    pk = PackedBatch(build_batch(per * rank, n_mol=per), pin=True)
    for _ in range(6):
        ts.step(pk, load=True)
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(K):
        ts.step(pk, load=False)
    b.record()
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b)], device=dev, dtype=torch.float64)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    ms = float(t[0]) / K
    return {"global_batch": 256, "molecules_per_rank": per, "ms_per_step": ms, "molecules_per_s": 256 / (ms * 1e-3),
            "scaling": "strong"}


def collate_bench(ts, dev, K):
    """the same training step, but every batch is a different set of molecules assembled ON THE
    DEVICE from a resident per-molecule store (store.MoleculeStore), against the host collate of the same batch."""
    import time
    import numpy as np
    from grappa_b200 import synthetic as S
    from grappa_b200.batching import PackedBatch
    from grappa_b200.graph import batch
    from grappa_b200.store import MoleculeStore
    mols = S.cfg2_graphs(3 * MOLS_PER_RANK, N_CONFS, first_seed=1000)
    store = MoleculeStore(mols, dev)
    rng = np.random.default_rng(0)
    draws = [rng.choice(len(mols), MOLS_PER_RANK, replace=False) for _ in range(K + 6)]
    for ids in draws[:6]:                              # warm-up incl. CUDA-graph capture for this batch shape
        ts.step_from_store(store, ids)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h2d = 0
    a.record()
    for ids in draws[6:]:
        loss = ts.step_from_store(store, ids)
    last = loss.item()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / K
    spec = store.prepare(draws[0])
    t0 = time.perf_counter()
    n_host = 5
    for ids in draws[:n_host]:
        PackedBatch(batch([mols[i] for i in ids]), pin=False)
    host_ms = (time.perf_counter() - t0) / n_host * 1e3
    t0 = time.perf_counter()
    for ids in draws[:n_host]:
        store.prepare(ids)
    prep_ms = (time.perf_counter() - t0) / n_host * 1e3
    return {"molecules_per_s": MOLS_PER_RANK / (ms * 1e-3), "ms_per_step": ms, "h2d_bytes_per_step": spec.h2d_bytes,
            "store_molecules": len(mols), "store_bytes": store.nbytes, "host_prepare_ms_per_batch": prep_ms,
            "host_collate_ms_per_batch_1_thread": host_ms, "final_loss": last,
            "note": "every step draws 32 different molecules; batch assembled by grappa_collate_float/int from the HBM-resident "
                    "store; the host uploads only the segment tables"}


def hbm_kernels(g, pk):
    """The gather / scatter kernels of the path timed alone at cfg-3 size (15 k atoms) against the HBM roofline:
    algorithmic bytes (K1 = (2N + E) * 2048 B; gathers = one read of the source rows + one write of the
    tokens) / device time from CUDA events over back-to-back launches.  Inputs exceed L2 or are rotated, no flush."""
    from grappa_b200 import ops
    from grappa_b200.plan import get_plan
    dev = g.nodes["n1"].data["partial_charge"].device
    plan = get_plan(g)
    N, E = plan.n_atoms, plan.n_edges
    res = {}

    def timed(fn, n=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / n

    with torch.no_grad():
        fts = [torch.randn(N, 512, device=dev) for _ in range(8)]          # 8 x 31 MB rotated: beyond L2 reuse
        i = [0]
        def k1():
            i[0] = (i[0] + 1) % len(fts)
            ops.gnn_attention(fts[i[0]], plan, 16)
        ms = timed(k1)
        byts = (2 * N + E) * 2048
        res["gnn_attn_fwd"] = {"ms": ms, "algorithmic_bytes": byts, "GBps": byts / ms / 1e6, "hbm_frac": byts / ms / 1e6 / pk["hbm_gbs"]}
        T4 = plan.T["n4"]
        names = ("tab_idx_n4", "tab_ptr_n4", "tab_tok_n4")
        tab = torch.randn(2 * N, 1536, device=dev)                          # the QKV table of the proper torsions
        ms = timed(lambda: ops.row_gather(tab, plan, names, 4 * T4))
        byts = (2 * N + 4 * T4) * 1536 * 4 + 4 * T4 * 4
        res["row_gather_fwd_qkv_n4"] = {"ms": ms, "algorithmic_bytes": byts, "GBps": byts / ms / 1e6, "hbm_frac": byts / ms / 1e6 / pk["hbm_gbs"]}
        qkv = torch.randn(4 * T4, 1536, device=dev)
        ms = timed(lambda: ops.tuple_attention(qkv, T4, 4, 8))
        byts = 4 * T4 * (1536 + 512) * 4
        res["tuple_attn_fwd_n4"] = {"ms": ms, "algorithmic_bytes": byts, "GBps": byts / ms / 1e6, "hbm_frac": byts / ms / 1e6 / pk["hbm_gbs"]}
        x = torch.randn(4 * T4, 512, device=dev)
        gam, bet = torch.ones(512, device=dev), torch.zeros(512, device=dev)
        ms = timed(lambda: ops.layer_norm(x, gam, bet))
        byts = 4 * T4 * 512 * 4 * 2
        res["layernorm_fwd_tokens_n4"] = {"ms": ms, "algorithmic_bytes": byts, "GBps": byts / ms / 1e6, "hbm_frac": byts / ms / 1e6 / pk["hbm_gbs"]}
    return res


def extras(dev, pk):
    """BASELINE.json configs[2] (inference atoms/s, 15k-atom system) and configs[3] (energy sweep, 1000 x 500)."""
    from grappa_b200 import synthetic as S
    from grappa_b200.energy import Energy
    from grappa_b200.model import GrappaModel
    out = {}
    torch.manual_seed(0)
    model = GrappaModel.from_default_config().to(dev).eval()
    g = S.cfg3_graph().to(dev)
    with torch.no_grad():
        for _ in range(3):
            model(g)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        a.record()
        for _ in range(n):
            model(g)
        b.record()
        torch.cuda.synchronize()
    ms = a.elapsed_time(b) / n
    out["inference_cfg3"] = {"atoms": g.num_nodes("n1"), "ms": ms, "atoms_per_s": g.num_nodes("n1") / (ms * 1e-3),
                             "algorithmic_tflop": 1.7, "tflops": 1.7 / (ms * 1e-3)}
    # The same system through the Grappa.predict boundary (grappa.py:62-73: g.to(device) -> model(g) -> results to the
    # host), one-shot: a HOST graph whose batch plan has never been built (cold plan: grappa_plan_build inside), the
    # host->device copies of features + plan, the forward, and the device->host read of every predicted parameter.
    # Wall clock around the whole call (it contains host work), device idle before and after.
    from grappa_b200.graph import HeteroGraph
    g_host = S.cfg3_graph()

    def fresh(gh):
        c = HeteroGraph(dict(gh._num_nodes), gh._edges, {k: v.clone() for k, v in gh._batch_num_nodes.items()})
        for nt in gh.ntypes:
            c.nodes[nt].data.update(gh.nodes[nt].data)
        return c

    def predict(gh):
        gd = gh.to(dev)
        with torch.no_grad():
            gd = model(gd)
        res = {}
        for lvl in ("n2", "n3", "n4", "n4_improper"):
            for key in ("k", "eq"):
                if key in gd.nodes[lvl].data:
                    res[(lvl, key)] = gd.nodes[lvl].data[key].to("cpu")
        return res

    for _ in range(2):
        predict(fresh(g_host))
    torch.cuda.synchronize()
    times = []
    for _ in range(5):
        gh = fresh(g_host)
        t0 = time.perf_counter()
        res = predict(gh)
        torch.cuda.synchronize()
        times.append((time.perf_counter() - t0) * 1e3)
    from grappa_b200.plan import build_plan
    t0 = time.perf_counter()
    cold_plan = build_plan(fresh(g_host))
    plan_ms = (time.perf_counter() - t0) * 1e3
    feat_bytes = sum(v.numel() * v.element_size() for nt in g_host.ntypes for v in g_host.nodes[nt].data.values())
    out["inference_cfg3_e2e"] = {"atoms": g_host.num_nodes("n1"), "ms_median": sorted(times)[len(times) // 2], "ms_all": times,
                                 "atoms_per_s": g_host.num_nodes("n1") / (sorted(times)[len(times) // 2] * 1e-3),
                                 "host_plan_ms": plan_ms, "h2d_bytes": int(feat_bytes + cold_plan.h2d_bytes),
                                 "d2h_bytes": int(sum(v.numel() * 4 for v in res.values())),
                                 "what": "host graph with no cached plan -> plan build (C++) -> H2D -> forward -> parameters to host"}
    out["hbm_kernels_cfg3"] = hbm_kernels(g, pk)
    del g
    g4 = S.cfg4_batch(1000, 500).to(dev)
    k_src = S.trimer_molecule(0, 1).to(dev)
    with torch.no_grad():
        k_src = model(k_src)
    for lvl in ("n2", "n3", "n4", "n4_improper"):
        for key in ("k", "eq"):
            if key in k_src.nodes[lvl].data:
                v = k_src.nodes[lvl].data[key]
                g4.nodes[lvl].data[key] = v.repeat(1000, *([1] * (v.dim() - 1))).contiguous()
    en = Energy(materialize_tuple_terms=False)
    with torch.no_grad():
        for _ in range(3):
            en(g4)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        a.record()
        for _ in range(n):
            en(g4)
        b.record()
        torch.cuda.synchronize()
    ms = a.elapsed_time(b) / n
    units = 1000 * 500
    gbs = units * (24 * 40 + 20) / (ms * 1e-3) / 1e9
    out["energy_cfg4"] = {"mol_conf_per_s": units / (ms * 1e-3), "ms": ms, "algorithmic_GBps": gbs, "hbm_frac": gbs / pk["hbm_gbs"]}
    return out


if __name__ == "__main__":
    main()
