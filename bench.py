#!/usr/bin/env python
"""bench.py -- DeepFM train samples/sec on Criteo-shaped synthetic data (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (N=1): configs[3] on one GPU: DeepFM, 13 cont + 26 single-valued cate, k=16, MLP 400-400-400,
batch 64K per GPU, ~1e8 table rows.  One "step" = one full train step (gather+FM, MLP fwd/bwd, log-loss,
sort/segment-reduce scatter, Adam).  Prints ONE JSON line (see the task contract).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "deepfm_train_samples_per_sec"
UNIT = "samples/s"


def hparams(batch, rows):
    from deep_learning_b200.hparams import ModelParams   # the reference's hyper-parameter bag (attribute names)
    return ModelParams(hidden_units=[400, 400, 400], batch_size=batch, embedding_size=16, cont_field_size=13,
                   cate_field_size=26, cate_feats_size=rows, l2_reg=1e-5, learning_rate=1e-3)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured"
    return 6650.0, 1400.0, "fallback"


class NvmlClockSampler:
    """SM clock / throttle reasons through NVML, every ~2 ms from a thread, plus one synchronous sample at each end of
    the timed region (the driver's 20-step region lasts ~20 ms: a 100 ms nvidia-smi poll cannot land inside it)."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index: int):
        import pynvml
        self.nv = pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:                                    # NVML counts physical devices; CUDA_VISIBLE_DEVICES may renumber them
            try:
                index = int(vis.split(",")[index])
            except ValueError:
                pass
        self.dev = pynvml.nvmlDeviceGetHandleByIndex(index)
        self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM))
        self.rows, self.on = [], False
        self.t = None

    def sample(self):
        nv = self.nv
        try:
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev)
        except Exception:
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
        self.rows.append((time.time(), float(nv.nvmlDeviceGetClockInfo(self.dev, nv.NVML_CLOCK_SM)), int(r)))

    def start(self):
        self.on = True
        self.sample()

        def loop():
            while self.on:
                self.sample()
                time.sleep(0.002)
        self.t = threading.Thread(target=loop, daemon=True)
        self.t.start()

    def stop(self, t0, t1):
        self.sample()
        self.on = False
        self.t.join(timeout=1.0)
        rows = [r for r in self.rows if t0 <= r[0] <= t1 + 0.005] or self.rows[-1:]
        sm = sorted(r[1] for r in rows)
        bits = 0
        for r in rows:
            bits |= r[2]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.mx, "reasons": [n for n, b in self.REASONS if bits & b],
                "samples": len(rows), "how": "NVML, sampled during the timed region"}


def clock_sampler(index: int):
    try:
        c = NvmlClockSampler(index)
        return c
    except Exception:
        return ClockSampler(index)


class ClockSampler:
    """fallback: nvidia-smi clocks / throttle reasons sampled DURING the timed region"""

    def start(self):
        pass

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            inside = t0 <= ts <= t1 + 0.1
            try:
                if inside:
                    sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            if inside:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_arm(steps, warmup, batch=4096, rows=2_000_000, adam_mode="tf_dense"):
    """The reference's CPU path, restated (oracle, fp32, all host threads): one train step =
    forward + log-loss + autodiff + TF1 dense Adam over the whole table, on a bounded sample of the
    workload (same field shapes; batch and table rows reduced so the run ends in tens of seconds)."""
    import torch
    from oracle import ctr_oracle as O
    import ctr_helpers as H
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    hp = hparams(batch, rows)
    p = O.init_params("deepfm", hp, dtype=torch.float32)
    st = O.adam_init(p)
    batches = [H.oracle_batch(H.make_batch("deepfm", hp, seed=2020 + i), torch.float32) for i in range(2)]
    for i in range(warmup):
        _, p = O.train_step("deepfm", p, st, batches[i % 2], hp, adam_mode=adam_mode)
    t0 = time.perf_counter()
    for i in range(steps):
        _, p = O.train_step("deepfm", p, st, batches[i % 2], hp, adam_mode=adam_mode)
    dt = time.perf_counter() - t0
    return {"value": batch * steps / dt, "unit": UNIT, "cores": cores, "kind": "port", "adam_mode": adam_mode,
            "sample": "oracle fp32 (PyTorch-CPU restatement of models/deepfm.py; TensorFlow 1.x not installable), "
                      "%d steps of batch %d, table %d rows, %s" % (
                          steps, batch, rows, "TF1 dense Adam (the reference's semantics: every row moves every step)"
                          if adam_mode == "tf_dense" else "lazy Adam (touched rows only: the GPU arm's headline mode)"),
            "ms_per_step": dt / steps * 1e3}


_JSON_OUT = None


def _guard_stdout():
    """stdout must carry exactly ONE JSON line, but libraries write to file descriptor 1 behind Python's back (NCCL prints
    its version banner and, at NCCL_DEBUG=INFO, its whole log there; pointing NCCL_DEBUG_FILE at /dev/stderr did not
    hold on the GPU box).  Keep a private duplicate of fd 1 for the line and make fd 1 an alias of stderr for everybody
    else: NCCL's log stays visible (the driver checks the communicator's rank count in it) and cannot corrupt the line."""
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line: dict) -> None:
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()


def main():
    _guard_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--batch", type=int, default=65536)
    ap.add_argument("--rows", type=int, default=100_000_000)
    ap.add_argument("--precision", default=os.environ.get("CTR_PRECISION", "tf32"))
    ap.add_argument("--dist", default="zipf")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--adam-mode", default="lazy", choices=("lazy", "tf_dense"))
    ap.add_argument("--no-extras", action="store_true", help="skip the untimed sub-records (fp32, tf_dense, sustained, TF32 peak)")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    W = max(a.warmup, 3)

    if a.impl == "reference":
        if rank != 0:
            return
        # exactly the K steps and W warm-ups asked for (one step of the sample is ~0.15 s on 16 cores); K capped at 200
        k_run, w_run = max(1, min(a.steps, 200)), max(1, min(a.warmup, 10))
        cb = cpu_reference_arm(k_run, w_run)
        line = {"impl": "reference", "adam_mode": "tf_dense", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": a.gpus,
                "steps": k_run, "warmup": w_run, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "DeepFM Criteo-shape 13 cont + 26 cate k=16 MLP 400x3 (CPU sample: " + cb["sample"] + ")"},
                "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return

    import torch
    import torch.distributed as dist
    from deep_learning_b200 import data as D
    from deep_learning_b200.engine import CtrEngine
    from deep_learning_b200.models._base import init_weights
    from deep_learning_b200.spec import spec_from_hparams

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    B, R = a.batch, a.rows
    hp = hparams(B, R)
    spec = spec_from_hparams("deepfm", hp)
    if world > 1:
        # NCCL_DEBUG is left as the launcher set it: its log goes to what is now stderr (_guard_stdout), not switched off
        # configs[3]: the ~1e8-row tables row-sharded over the ranks (owner = id mod N), B samples per GPU (weak scaling)
        dist.init_process_group("nccl", device_id=dev)
        from deep_learning_b200.sharded import ShardedCtrEngine
        eng = ShardedCtrEngine(spec, B, device=dev, learning_rate=hp.learning_rate, precision=a.precision, adam_mode=a.adam_mode)
    else:
        eng = CtrEngine(spec, B, device=dev, learning_rate=hp.learning_rate, precision=a.precision, adam_mode=a.adam_mode)
    eng.init_table_on_device(seed=2019)
    eng.set_dense_weights(init_weights(spec, 2019, with_table=False))     # same seed on every rank: replicas start equal

    # rotating synthetic batches, generated on the device (seed 2020 + i*world + rank), mirrored in pinned host memory
    sizes = D.field_sizes(R, spec.Fc)
    NB = 4
    dev_batches = [D.synthetic_batch(B, spec.C, spec.Fc, sizes, seed=2020 + i * world + rank, dist=a.dist, device=dev)
                   for i in range(NB)]
    host_batches = [{k: v.cpu().pin_memory() for k, v in b.items()} for b in dev_batches]
    h2d = sum(v.numel() * v.element_size() for k, v in host_batches[0].items() if k != "vector_feats")

    def sync():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident timing ------------------------------------------------------------------
    for i in range(W):
        eng.train_step(dev_batches[i % NB], next_batch=dev_batches[(i + 1) % NB])
    sync()
    eng.h.timers = {"ctr_fm_gather_fwd": []}      # the roofline kernel is timed inside the timed region (one event pair per step)
    l0 = eng.h.launches
    clocks = clock_sampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    clocks.start()
    t0 = time.time()
    torch.cuda.profiler.start()      # `ncu --profile-from-start off` captures exactly the timed steps
    # behind the host barrier + synchronize the ranks are still milliseconds apart (8 processes leaving an NCCL barrier):
    # one device-side rank barrier puts every rank's start event at the same instant, so a short timed window measures
    # steps and not the start-up skew (it was ~0.1 ms per step over 30 steps at N=8: 1.36 ms against 1.25 per rank)
    if world > 1:
        eng.align_ranks()
    e0.record()
    for i in range(a.steps):
        eng.train_step(dev_batches[i % NB], next_batch=dev_batches[(i + 1) % NB])
    e1.record()
    sync()
    torch.cuda.profiler.stop()
    t1 = time.time()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = eng.h.launches - l0
    k1_ev = eng.h.timers["ctr_fm_gather_fwd"]
    k1_ms = sum(x.elapsed_time(y) for x, y in k1_ev) / max(len(k1_ev), 1)
    clk = clocks.stop(t0, t1)
    # ---- per-kernel table: the same steps again with a CUDA event pair around each selected entry point, on the stream
    # it launches on.  Kept out of the timed region above (only the roofline kernel is timed there): ~60 extra event
    # records per step are host work, and the multi-GPU step is close to host-bound.
    timed = ("ctr_fm_gather_fwd", "ctr_embed_bwd_partials_weighted", "ctr_adam_rows", "ctr_adam_rows_gathered",
             "ctr_sort_ids", "ctr_unique_segments", "ctr_mlp_fwd", "ctr_mlp_fwd_bits", "ctr_mlp_bwd_data", "ctr_mlp_bwd_data_bits",
             "ctr_mlp_bwd_weight", "ctr_head_fwd",
             "ctr_fm_cont_bwd", "ctr_dense_update", "ctr_gather_rows_push", "ctr_embed_bwd_finalize_push", "ctr_peer_barrier")
    eng.h.timers = {n: [] for n in timed}
    for i in range(a.steps):
        eng.train_step(dev_batches[i % NB], next_batch=dev_batches[(i + 1) % NB])
    sync()
    kern = {n: (sum(x.elapsed_time(y) for x, y in ev) / max(len(ev), 1), len(ev)) for n, ev in eng.h.timers.items()}
    eng.h.timers = None
    n_unique = int(eng.n_unique.item())
    value = world * B * a.steps / (ms * 1e-3)

    # ---- end to end through the public step API: pinned host batch in, loss scalar out, every step ----
    for i in range(2):
        float(eng.train_step(host_batches[i % NB], next_batch=host_batches[(i + 1) % NB]).item())
    sync()
    if world > 1:
        eng.align_ranks()
    e0.record()
    for i in range(a.steps):
        eng.train_step(host_batches[i % NB], next_batch=host_batches[(i + 1) % NB])
        loss = eng.last_loss()       # this step's loss on the host, every step (waits for its D2H copy, see CtrEngine.last_loss)
    e1.record()
    sync()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1))

    layer_dims, table_gb = eng._layer_dims(), eng.table.numel() * 4 / 1e9
    exch = ({"unique_rows_requested_rank0": int(eng.U), "rows_served_rank0": int(eng.n_req),
             "wire_bytes_per_row": 4 + 2 * 4 * eng.wire} if world > 1 else None)

    # ---- untimed sub-records (one GPU only): the regimes BASELINE.md section 4 lists beside the headline --------------
    extras = {}
    if world == 1 and not a.no_extras:
        def timed_steps(engine, batches, steps, warm=3, k1=False):
            for i in range(warm):
                engine.train_step(batches[i % len(batches)], next_batch=batches[(i + 1) % len(batches)])
            sync()
            engine.h.timers = {"ctr_fm_gather_fwd": []} if k1 else None
            x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            x0.record()
            for i in range(steps):
                engine.train_step(batches[i % len(batches)], next_batch=batches[(i + 1) % len(batches)])
            x1.record()
            sync()
            ev = engine.h.timers["ctr_fm_gather_fwd"] if k1 else []
            engine.h.timers = None
            t = x0.elapsed_time(x1) / steps
            return t, (sum(x.elapsed_time(y) for x, y in ev) / len(ev) if ev else None)

        # (a) the same loop for >= 3 s: clocks settle under the power cap (the 20-step region above runs at burst clocks)
        n_sus = max(int(3000.0 / (ms / a.steps)), a.steps)
        cs = clock_sampler(local)
        cs.start()
        ts0 = time.time()
        t_sus, _ = timed_steps(eng, dev_batches, n_sus, warm=0)
        extras["sustained"] = {"value": B / (t_sus * 1e-3), "unit": UNIT, "ms_per_step": t_sus, "steps": n_sus,
                               "clocks": cs.stop(ts0, time.time())}
        # (b) uniform ids: no duplicates inside a batch, every row read misses L2 -- the regime where K1 is an HBM test
        ub = [D.synthetic_batch(B, spec.C, spec.Fc, sizes, seed=3020 + i, dist="uniform", device=dev) for i in range(NB)]
        t_u, k1_u = timed_steps(eng, ub, 10, k1=True)
        extras["uniform_ids"] = {"value": B / (t_u * 1e-3), "unit": UNIT, "ms_per_step": t_u, "k1_ms": k1_u,
                                 "unique_rows_last_step": int(eng.n_unique.item())}
        del ub
        # (c) TF32 GEMM rate of the library GEMM on this GPU (cuBLAS through torch.matmul, allow_tf32), burst: the
        # denominator of mlp.frac -- MEASURED_PEAKS.json holds bf16 only
        torch.backends.cuda.matmul.allow_tf32 = True
        ga, gb = torch.randn(8192, 8192, device=dev), torch.randn(8192, 8192, device=dev)
        best = 1e9
        for i in range(8):
            x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            x0.record()
            gc = ga @ gb
            x1.record()
            torch.cuda.synchronize(dev)
            if i >= 2:
                best = min(best, x0.elapsed_time(x1))
        extras["tf32_gemm_peak_tflops"] = 2 * 8192 ** 3 / (best * 1e-3) / 1e12
        del ga, gb, gc
    if world == 1 and not a.no_extras:
        # (d) the reference's arithmetic type and (e) its optimizer semantics, on the same workload
        host_w = init_weights(spec, 2019, with_table=False)
        del eng
        torch.cuda.empty_cache()
        for key, prec, mode, steps in (("fp32", "fp32", a.adam_mode, 10), ("tf_dense", a.precision, "tf_dense", 5)):
            e2 = CtrEngine(spec, B, device=dev, learning_rate=hp.learning_rate, precision=prec, adam_mode=mode)
            e2.init_table_on_device(seed=2019)
            e2.set_dense_weights(host_w)
            t2, _ = timed_steps(e2, dev_batches, steps, warm=3)
            extras[key] = {"value": B / (t2 * 1e-3), "unit": UNIT, "ms_per_step": t2, "steps": steps,
                           "mlp_precision": prec, "adam_mode": mode}
            del e2
            torch.cuda.empty_cache()
        eng = None
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the embedding gather (K1): algorithmic bytes/sample x B / measured duration ----
    hbm, tfl, which = peaks()
    k1_bytes = (4 * spec.Fc + 4 * spec.C + 4 * spec.k * spec.Fc + 4 * spec.Fc + 8) * B      # 1932 B/sample
    k1_gbs = k1_bytes / (k1_ms * 1e-3) / 1e9
    # DRAM bytes of one K1 launch from the committed `ncu --set full` capture of this command (profiles/): the only
    # number in this line that is not measured in this run
    traffic = None
    try:
        if world == 1 and B == 65536:
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "k1_ncu_traffic.json")) as f:
                tj = json.load(f)
            traffic = float(tj["dram_bytes_read"]) + float(tj["dram_bytes_write"])
    except Exception:
        traffic = None
    kernels = {}
    step_ms = ms / a.steps
    for n, (avg, cnt) in kern.items():
        per_step = avg * cnt / a.steps
        kernels[n] = {"ms_per_call": round(avg, 4), "calls_per_step": cnt / a.steps, "share_of_step": round(per_step / step_ms, 4)}
    mlp_flops = 3 * 2 * sum(k_ * h_ for k_, h_ in layer_dims) * B
    mlp_ms = sum(kern[n][0] * kern[n][1] for n in ("ctr_mlp_fwd", "ctr_mlp_fwd_bits", "ctr_mlp_bwd_data", "ctr_mlp_bwd_data_bits",
                                                     "ctr_mlp_bwd_weight")) / a.steps
    tf32_peak = extras.get("tf32_gemm_peak_tflops")
    mlp_tfl = mlp_flops / (mlp_ms * 1e-3) / 1e12 if mlp_ms > 0 else None
    line = {
        "metric": METRIC, "adam_mode": a.adam_mode, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": W,
        "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if a.precision == "fp32" else "tf32", "data": "synthetic",
        "config": {"workload": "DeepFM Criteo-shape: 13 cont + 26 single-valued cate, k=16, MLP 400-400-400, "
                               "batch %d per GPU, table %d rows x (16+1) f32%s, ids %s, Adam %s"
                               % (B, R, " row-sharded over %d GPUs (ids / vectors / gradients / dense-gradient sum exchanged through NVLink peer memory, "
                                  "no collective in the step)" % world
                                  if world > 1 else "", a.dist,
                                  "lazy (touched rows only; not the reference's semantics, see tf_dense sub-record)"
                                  if a.adam_mode == "lazy" else "tf_dense (TF1 semantics: every row moves every step)"),
                   "batch_per_gpu": B, "global_batch": B * world, "parallelism": "dp%d+rowshard%d" % (world, world) if world > 1 else "single",
                   "table_rows": R, "unique_rows_last_step": n_unique, "mlp_precision": a.precision,
                   "l2_flush": "working set per step (~0.9 GB activations + random rows of a %.1f GB table) >> 126 MB L2; "
                               "%d rotating batches" % (table_gb, NB)},
        "clocks": clk,
        "e2e": {"value": world * B * a.steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": 4 * world},
        "gpu_launches": int(launches),
        "roofline": {"kernel": "fm_gather_fast_kernel<16,true,false> (ctr_fm_gather_fwd)", "bound": "hbm", "achieved": k1_gbs, "peak": hbm,
                     "unit": "GB/s", "frac": k1_gbs / hbm, "traffic": traffic, "peak_source": which,
                     "algorithmic_bytes_per_launch": k1_bytes},
        "mlp": {"achieved_tflops": mlp_tfl, "ms_per_step": mlp_ms, "bound": "tensor (TF32) / HBM for the fp32 activations",
                "peak_tflops": tf32_peak, "peak_source": "cuBLAS TF32 8192^3 through torch.matmul, measured in this run (burst)",
                "frac": (mlp_tfl / tf32_peak) if (mlp_tfl and tf32_peak) else None},
        "kernels": kernels,
    }
    for k in ("sustained", "uniform_ids", "fp32", "tf_dense"):
        if k in extras:
            line[k] = extras[k]
    if "uniform_ids" in extras and extras["uniform_ids"]["k1_ms"]:
        u = extras["uniform_ids"]
        u["k1_achieved_gbs"] = k1_bytes / (u["k1_ms"] * 1e-3) / 1e9
        u["k1_frac"] = u["k1_achieved_gbs"] / hbm
    if exch is not None:
        line["exchange"] = exch
    if not a.no_cpu_baseline and world == 1:
        cb = cpu_reference_arm(60, 2)          # ~10 s of CPU work on the box's host cores
        line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "adam_mode")}
        cl = cpu_reference_arm(60, 2, adam_mode="lazy")      # like for like with the GPU arm's lazy Adam
        line["cpu_baseline_lazy"] = {k: cl[k] for k in ("value", "unit", "cores", "kind", "sample", "adam_mode")}
    emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
