#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 acquisition engine (contract in the task brief).

Headline workload (BASELINE.json configs[3], the one the metric is quoted on): max_EI candidate screening, km
matern5_2, n = 2000, d = 10, over 10^7 uniform candidates IN TOTAL.  With --gpus N the 10^7 candidates are cut into N
contiguous ranges (STRONG scaling: total work fixed, the model is replicated, the only collective is the all-gather of
each rank's (value, index) pair for the global argmax); the candidate matrix is the same whatever N (generated in 10^6-row
blocks from a counter-based seed), so the global argmax must not move with N.

  value : candidates/s, candidates already resident in HBM when the timed region starts
          (dk_screen_ei with a device pointer: fused EI kernel + device which.max + 16 B result).
  e2e   : the same call with HOST buffers in PAGEABLE memory (what an R vector is): the library streams them through the
          device in chunks (pinned staging, copies overlapped with the kernels); H2D of every candidate and D2H of the
          result are inside the timed region.  extra.e2e_pinned is the same with a pinned caller buffer.
  roofline : FP64 tensor pipe; achieved = algorithmic flops (n^2 + 4n + 2np per candidate) / CUDA-event time
          of the fused solve kernel; peak = cuBLAS DGEMM 8192^3 measured IN THIS RUN on the same GPU (MEASURED_PEAKS.json
          has no FP64 entry; the round-1 measurement is profiles/dgemm_peak_r01.json).
  cpu_baseline : the oracle port (numpy + scipy/OpenBLAS dtrsm) on a bounded candidate sample, all host threads.

Secondary lines under "extra", measured on EVERY rank over that rank's share and summed (the metric also names them):
  ei_grad : EI + EI.grad on cfg4 (2^21 candidates in total), device-resident value, e2e, roofline, cpu_baseline
  qei     : qEI(q = 4) on cfg3 (rosenbrock4, n = 500, 10^5 batches in total), same keys; roofline for both kernels
  cfg5    : AKG + kriging.quantile, d = 20, n = 5000, noise.var and nugget variants; the full 10^7 candidates when
            --gpus 8 (the named configuration) or --cfg5-full, else 2^16 per GPU
  weak    : the round-1 weak-scaling figure (10^7 candidates per GPU), only with --weak
  single_process : dk_multi_screen_ei over all N GPUs from ONE process (rank 0) with the whole host matrix

`--impl reference` times the reference's CPU algorithm (the oracle port: R is not installed here or on the GPU box,
profiles/probe_rscript_r02.txt; the arithmetic lives in DiceKriging) on the host cores for the same metric and config.
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
sys.path.insert(0, ROOT)

CFG = "cfg4_sphere10_n2000"
CFG3 = "cfg3_rosenbrock4_n500"
CFG5 = "cfg5_noisy20_n5000"
METRIC = "EI max_EI-screen candidates/sec at n=2000,d=10 (km matern5_2)"
UNIT = "candidates/s"
BLOCK = 1_000_000   # rows per block of the counter-based candidate generator
QN01 = -1.2815515655446004


def flops_per_candidate(n, p):
    return n * n + 4 * n + 2 * n * p


def candidate_rows(cfg, lo, hi, d, seed_extra=0):
    """Rows [lo, hi) of the global candidate matrix, as a (d, hi - lo) C-ordered array (= column-major rows x d).
    Block b of 10^6 rows comes from default_rng([cand_seed, seed_extra, b]): any rank can build any range."""
    import numpy as np
    from diceoptim_b200 import synth
    seed = synth.CONFIGS[cfg]["cand_seed"]
    out = np.empty((d, hi - lo))
    b0, b1 = lo // BLOCK, (hi + BLOCK - 1) // BLOCK
    for b in range(b0, b1):
        blk = np.random.default_rng([seed, seed_extra, b]).random((d, BLOCK))
        s, e = max(lo, b * BLOCK), min(hi, (b + 1) * BLOCK)
        out[:, s - lo:e - lo] = blk[:, s - b * BLOCK:e - b * BLOCK]
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (profiling recipe's clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [t.strip() for t in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------
# CPU legs (the oracle is the checker / the reported baseline, never the product)
# ---------------------------------------------------------------------------------------------------------
def cpu_oracle_rate(km, x, chunk=1000):
    """Oracle EI + which.max over the rows of x on ALL host cores: candidate chunks are spread over a thread pool
    (numpy and LAPACK release the GIL), BLAS kept single-threaded inside each worker; returns (seconds, argmax)."""
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    from threadpoolctl import threadpool_limits
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import to_oracle
    from oracle import diceoptim as do
    ko = to_oracle(km)
    parts = [x[i:i + chunk] for i in range(0, x.shape[0], chunk)]
    t0 = time.perf_counter()
    with threadpool_limits(limits=1), ThreadPoolExecutor(max_workers=os.cpu_count()) as pool:
        ei = np.concatenate(list(pool.map(lambda p: do.EI(p, ko), parts)))
    imax = do.which_max(ei)
    return time.perf_counter() - t0, imax


def cpu_ei_grad_rate(km, x):
    """Oracle EI.grad (one point per call, as the reference) over the rows of x, threads over points."""
    from concurrent.futures import ThreadPoolExecutor
    from threadpoolctl import threadpool_limits
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import to_oracle
    from oracle import diceoptim as do
    ko = to_oracle(km)
    t0 = time.perf_counter()
    with threadpool_limits(limits=1), ThreadPoolExecutor(max_workers=os.cpu_count()) as pool:
        list(pool.map(lambda xi: do.EI_grad(xi, ko), list(x)))
    return time.perf_counter() - t0


_QW = {}


def _qei_init(cfg):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import to_oracle
    from diceoptim_b200 import synth
    _QW["km"] = to_oracle(synth.config_model(cfg))


def _qei_one(args):
    from oracle import diceoptim as do
    xb, plugin = args
    return do.qEI(xb, _QW["km"], plugin=plugin, fastCompute=False)


def cpu_qei_rate(cfg, xb, plugin):
    """Oracle qEI(fastCompute = FALSE) (R/qEI.R:119-194) over the batches xb, one worker PROCESS per host core (the
    CDF quadrature is pure Python); returns (seconds, values).  Pool start-up is outside the timed region."""
    import multiprocessing as mpx
    workers = os.cpu_count() or 1
    with mpx.get_context("spawn").Pool(workers, initializer=_qei_init, initargs=(cfg,)) as pool:
        pool.map(_qei_one, [(xb[0], plugin)] * workers)          # warm every worker
        t0 = time.perf_counter()
        vals = pool.map(_qei_one, [(b, plugin) for b in xb], chunksize=max(1, len(xb) // (4 * workers)))
        dt = time.perf_counter() - t0
    return dt, vals


def run_reference(args):
    """The reference's CPU algorithm for the path (oracle port) on all host threads; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    from diceoptim_b200 import synth
    km = synth.config_model(CFG)
    sample = args.cpu_sample
    x = np.ascontiguousarray(candidate_rows(CFG, 0, sample, km.d).T)
    for _ in range(min(args.warmup, 1)):
        cpu_oracle_rate(km, x[:2000])
    times = []
    for _ in range(args.steps):
        dt, _ = cpu_oracle_rate(km, x)
        times.append(dt)
    total = sum(times)
    value = sample * args.steps / total
    cores = os.cpu_count()
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": CFG, "n": km.n, "d": km.d, "kernel": "matern5_2", "candidates_per_step": sample,
                   "note": "bounded sample of the 10^7-candidate workload; the rate is per candidate"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{sample} candidates/step x {args.steps} steps, numpy + scipy OpenBLAS dtrsm, chunks over {cores} threads"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--candidates", type=int, default=10_000_000, help="candidates per step IN TOTAL (strong scaling)")
    ap.add_argument("--cpu-sample", type=int, default=64000, help="candidates in the CPU-baseline sample")
    ap.add_argument("--no-extras", action="store_true", help="headline line only")
    ap.add_argument("--weak", action="store_true", help="also time 10^7 candidates PER GPU (round-1 weak figure)")
    ap.add_argument("--cfg5-full", action="store_true", help="cfg5 over the full 10^7 candidates whatever --gpus")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)

    import numpy as np
    import torch
    import torch.distributed as dist
    import ctypes as C
    import __graft_entry__ as g
    g.build()
    from diceoptim_b200 import synth, _lib, parallel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    host_group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        host_group = dist.new_group(backend="gloo")     # CPU-side barrier: ranks that wait must not spin a kernel on their GPU
    lib = _lib.load()
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sptr = C.c_void_p(stream.cuda_stream)
    vp = lambda t: C.c_void_p(t.data_ptr() if hasattr(t, "data_ptr") else t.ctypes.data)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allsum(v):
        if world == 1:
            return float(v)
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def allmax(v):
        if world == 1:
            return float(v)
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(fn, steps, sampler=None):
        """K calls of fn bracketed by barrier + synchronize, CUDA events on the launch stream, max over ranks (ms)."""
        barrier()
        if sampler:
            sampler.start()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        out = None
        for _ in range(steps):
            out = fn()
        e1.record(stream)
        barrier()
        return allmax(e0.elapsed_time(e1)), out

    def dgemm_peak():
        """cuBLAS DGEMM 8192^3, best of 10, on this GPU in this run: the FP64 tensor roofline denominator."""
        n = 8192
        a = torch.randn(n, n, dtype=torch.float64, device="cuda"); b = torch.randn(n, n, dtype=torch.float64, device="cuda")
        c = torch.empty_like(a)
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(10):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(stream); torch.matmul(a, b, out=c); e1.record(stream); e1.synchronize()
            best = min(best, e0.elapsed_time(e1))
        del a, b, c
        torch.cuda.empty_cache()
        return 2 * n ** 3 / best * 1e-9

    # ------------------------------------------------------------------ headline: cfg4 screening, strong scaling
    km = synth.config_model(CFG)
    n, d, p = km.n, km.d, km.p
    h = km.handle(local)
    lib.dk_model_set_timing(h.ptr, 1)
    NT = args.candidates
    lo, hi = parallel.shard_range(NT, rank, world)
    N = hi - lo
    plugin = float(np.min(km.y))
    x_page = candidate_rows(CFG, lo, hi, d)                # pageable host memory (numpy), column-major N x d
    x_dev = torch.from_numpy(x_page).to("cuda")
    idx = np.empty(1, dtype=np.int64); val = np.empty(1)

    def screen(ptr, on_device, count=N):
        _lib.check(lib.dk_screen_ei(h.ptr, ptr, 1 if on_device else 0, count, plugin, 1, 1, 0, 1,
                                    vp(idx), vp(val), sptr if on_device else None), "dk_screen_ei")
        v, gi = parallel.global_argmax(float(val[0]), int(idx[0]), lo, device="cuda")
        return gi, v

    kernel_ms = []

    def step_dev():
        r = screen(vp(x_dev), True)
        kernel_ms.append(lib.dk_model_last_kernel_ms(h.ptr))
        return r

    for _ in range(args.warmup):
        step_dev()
    kernel_ms.clear()
    l0 = h.launches()
    sampler = ClockSampler(local)
    ms_dev, res_dev = timed(step_dev, args.steps, sampler)
    clocks = sampler.stop()
    launches = h.launches() - l0
    step_page = lambda: screen(vp(x_page), False)
    step_page()
    ms_e2e, res_e2e = timed(step_page, args.steps)
    assert res_dev == res_e2e, (res_dev, res_e2e)
    value = NT * args.steps / (ms_dev * 1e-3)
    e2e = NT * args.steps / (ms_e2e * 1e-3)

    peak = dgemm_peak()
    try:
        peak_r01 = float(json.load(open(os.path.join(ROOT, "profiles", "dgemm_peak_r01.json")))["burst_tflops"])
    except Exception as exc:
        raise SystemExit(f"profiles/dgemm_peak_r01.json missing or unreadable ({exc}): the roofline has no committed reference")
    peak_src = "cuBLAS DGEMM 8192^3 (torch.matmul f64), best of 10, measured in this run on this GPU"
    k_ms = sum(kernel_ms) / len(kernel_ms)
    achieved = flops_per_candidate(n, p) * N / (k_ms * 1e-3) * 1e-12
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
        traffic = tj["k_predict_fused_bytes_per_candidate"] * N
    except Exception:
        pass

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": CFG, "n": n, "d": d, "p": p, "kernel": "matern5_2", "type": "UK",
                   "candidates_total_per_step": NT, "candidates_this_rank": N,
                   "parallelism": f"10^7 candidates cut into {world} contiguous ranges, model replicated",
                   "l2": "inputs (80 B/candidate, >= 100 MB per rank) and per-CTA panels (296 MB) exceed the 126 MB L2; no flush needed",
                   "argmax": {"index": res_dev[0], "value": res_dev[1]}},
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(N * d * 8), "d2h_bytes_per_step": 16,
                "ms_per_step": ms_e2e / args.steps, "host_memory": "pageable (numpy), streamed in chunks through pinned staging"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                     "traffic": traffic, "kernel": "k_predict_fused_tma", "kernel_ms": k_ms, "candidates_per_launch": N,
                     "flops_per_candidate": flops_per_candidate(n, p), "peak_source": peak_src,
                     "peak_r01_profile": peak_r01},
        "clocks": clocks,
    }
    extra = {}

    if not args.no_extras:
        # ---- pinned caller buffer (round 1's e2e) ----
        x_pin = torch.empty((d, N), dtype=torch.float64, pin_memory=True)
        x_pin.numpy()[:] = x_page
        step_pin = lambda: screen(vp(x_pin), False)
        step_pin()
        ms_pin, res_pin = timed(step_pin, args.steps)
        assert res_pin == res_dev
        extra["e2e_pinned"] = {"value": NT * args.steps / (ms_pin * 1e-3), "unit": UNIT, "ms_per_step": ms_pin / args.steps}
        del x_pin

        # ---- the same screen with the candidates GENERATED on the device (dk_screen + dk_uniform_spec): no upload at all ----
        gen = _lib.dk_uniform_spec(seed=1000 + rank, lower=None, upper=None)
        crit = _lib.dk_crit_spec(criterion=0, type=1, minimization=1, proxy=0, plugin=plugin, qnorm_beta=0.0, new_noise_var=0.0)
        gi = np.empty(1, dtype=np.int64); gv = np.empty(1); gx = np.empty((1, d), order="F")
        step_gen = lambda: _lib.check(lib.dk_screen(h.ptr, None, 0, C.byref(gen), N, C.byref(crit), 1, vp(gi), vp(gv), vp(gx), None), "dk_screen")
        step_gen()
        ms_gen, _ = timed(step_gen, args.steps)
        extra["e2e_generated"] = {"value": NT * args.steps / (ms_gen * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 0,
                                  "d2h_bytes_per_step": 16, "note": "uniform candidates drawn on the device by the counter-based generator "
                                  "(what lower + runif(d N) (upper - lower) of R/max_crit.R:104 draws on the host); every rank its own seed"}

        # ---- EI + EI.grad on cfg4, every rank on its share of 2^21 candidates ----
        NG = 1 << 21
        glo, ghi = parallel.shard_range(NG, rank, world)
        Ng = ghi - glo
        xg_host = np.ascontiguousarray(x_page[:, :Ng])
        xg = x_dev[:, :Ng].contiguous()
        grad = torch.empty((d, Ng), dtype=torch.float64, device="cuda")
        eiv = torch.empty(Ng, dtype=torch.float64, device="cuda")
        grad_h = np.empty((d, Ng)); ei_h = np.empty(Ng)
        g_ms = []

        def gstep():
            _lib.check(lib.dk_ei_grad_dev(h.ptr, vp(xg), Ng, plugin, 1, 1, 0, vp(grad), vp(eiv), sptr), "dk_ei_grad_dev")
            g_ms.append(lib.dk_model_last_kernel_ms(h.ptr))

        def gstep_host():
            _lib.check(lib.dk_ei_grad(h.ptr, vp(xg_host), Ng, plugin, 1, 1, 0, vp(grad_h), vp(ei_h)), "dk_ei_grad")

        gstep(); gstep(); g_ms.clear()
        gms, _ = timed(gstep, args.steps)
        gstep_host()
        gms_h, _ = timed(gstep_host, args.steps)
        gfl = 2 * n * n + 2 * n * d + 4 * n
        gk = sum(g_ms) / len(g_ms)
        g_ach = allsum(gfl * Ng / (gk * 1e-3) * 1e-12) / world      # per-GPU average
        extra["ei_grad"] = {
            "metric": "EI + EI.grad candidates/s at n=2000,d=10", "value": NG * args.steps / (gms * 1e-3), "unit": UNIT,
            "candidates_total_per_step": NG, "scaling": "strong",
            "e2e": {"value": NG * args.steps / (gms_h * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(Ng * d * 8),
                    "d2h_bytes_per_step": int(Ng * (d + 1) * 8), "host_memory": "pageable"},
            "roofline": {"bound": "tensor", "achieved": g_ach, "peak": peak, "unit": "TFLOP/s", "frac": g_ach / peak,
                         "kernel": "k_grad_fused", "kernel_ms": gk, "flops_per_candidate": gfl, "per_gpu": True},
        }
        del grad, eiv, xg

        # ---- qEI(q = 4) on cfg3, every rank on its share of 10^5 batches ----
        km3 = synth.config_model(CFG3)
        h3 = km3.handle(local)
        lib.dk_model_set_timing(h3.ptr, 1)
        BT, q = 100_000, 4
        blo, bhi = parallel.shard_range(BT, rank, world)
        Bq = bhi - blo
        xq_host = candidate_rows(CFG3, blo * q, bhi * q, km3.d)          # rows b q .. b q + q - 1 form batch b
        xq = torch.from_numpy(xq_host).to("cuda")
        oq = torch.empty(Bq, dtype=torch.float64, device="cuda")
        oq_h = np.empty(Bq)
        pl3 = float(np.median(km3.y))
        q_ms, q_ms_first = [], []

        def qstep():
            _lib.check(lib.dk_qei_dev(h3.ptr, vp(xq), Bq, q, pl3, 1, 1, vp(oq), sptr), "dk_qei_dev")
            q_ms.append(lib.dk_model_last_kernel_ms(h3.ptr))
            q_ms_first.append(lib.dk_model_last_phase_ms(h3.ptr))

        def qstep_host():
            _lib.check(lib.dk_qei(h3.ptr, vp(xq_host), Bq, q, pl3, 1, 1, vp(oq_h)), "dk_qei")

        qstep(); qstep(); q_ms.clear(); q_ms_first.clear()
        qms, _ = timed(qstep, max(args.steps, 5))
        qstep_host()
        qms_h, _ = timed(qstep_host, max(args.steps, 5))
        qsteps = max(args.steps, 5)
        mom_ms = sum(q_ms_first) / len(q_ms_first); both_ms = sum(q_ms) / len(q_ms)
        mom_fl = q * km3.n ** 2 + q * q * km3.n
        mom_ach = allsum(mom_fl * Bq / (mom_ms * 1e-3) * 1e-12) / world
        cdf_fp64 = json.load(open(os.path.join(ROOT, "profiles", "qei_cdf_work.json")))
        cdf_ach = allsum(cdf_fp64["fp64_flops_per_eval_q4"] * Bq / ((both_ms - mom_ms) * 1e-3) * 1e-12) / world
        assert np.array_equal(oq.cpu().numpy(), oq_h, equal_nan=True)
        extra["qei"] = {
            "metric": "qEI(q=4) evals/s at cfg3 (rosenbrock4 km, n=500, d=4)", "value": BT * qsteps / (qms * 1e-3),
            "unit": "evals/s", "batches_total_per_step": BT, "scaling": "strong", "nan": int(torch.isnan(oq).sum().item()),
            "e2e": {"value": BT * qsteps / (qms_h * 1e-3), "unit": "evals/s", "h2d_bytes_per_step": int(Bq * q * km3.d * 8),
                    "d2h_bytes_per_step": int(Bq * 8), "host_memory": "pageable"},
            "roofline": [
                {"kernel": "k_qei_moments", "bound": "tensor", "achieved": mom_ach, "peak": peak, "unit": "TFLOP/s",
                 "frac": mom_ach / peak, "kernel_ms": mom_ms, "flops_per_eval": mom_fl, "per_gpu": True},
                {"kernel": "k_qei_cdf", "bound": "fp64 pipe", "achieved": cdf_ach, "peak": cdf_fp64["fp64_pipe_peak_tflops"],
                 "unit": "TFLOP/s", "frac": cdf_ach / cdf_fp64["fp64_pipe_peak_tflops"], "kernel_ms": both_ms - mom_ms,
                 "flops_per_eval": cdf_fp64["fp64_flops_per_eval_q4"], "source": cdf_fp64["source"], "per_gpu": True}],
        }

        # ---- cfg5: AKG + kriging.quantile, n = 5000, d = 20, noise.var and nugget ----
        full5 = args.cfg5_full or world == 8
        # below the full configuration: a sample of four whole waves of the persistent grid per GPU (2 CTAs per SM x 64
        # candidates), so that the rate is not that of a half-empty last wave (the full 10^7 is 66 waves per GPU)
        wave = 2 * torch.cuda.get_device_properties(local).multi_processor_count * 64
        N5T = 10_000_000 if full5 else 4 * wave * world
        a_lo, a_hi = parallel.shard_range(N5T, rank, world)
        N5 = a_hi - a_lo
        x5_host = candidate_rows(CFG5, a_lo, a_hi, 20)
        x5 = torch.from_numpy(x5_host).to("cuda")
        o5 = torch.empty(N5, dtype=torch.float64, device="cuda")
        o5_h = np.empty(N5)
        extra["cfg5"] = {"candidates_total_per_step": N5T, "full_config": bool(full5)}
        for kind in ("noise_var", "nugget"):
            km5 = synth.config_model(CFG5, noise_kind=kind)
            h5 = km5.handle(local)
            lib.dk_model_set_timing(h5.ptr, 1)
            res = {}
            calls = {
                "akg": (lambda: lib.dk_akg_dev(h5.ptr, vp(x5), N5, 0.1, 1, vp(o5), sptr),
                        lambda: lib.dk_akg(h5.ptr, vp(x5_host), N5, 0.1, 1, vp(o5_h)), 2 * km5.n ** 2),
                "quantile": (lambda: lib.dk_quantile_dev(h5.ptr, vp(x5), N5, QN01, 1, vp(o5), None, sptr),
                             lambda: lib.dk_quantile(h5.ptr, vp(x5_host), N5, QN01, 1, vp(o5_h), None), km5.n ** 2),
            }
            for name, (dev_call, host_call, fl) in calls.items():
                k5 = []

                def dstep():
                    _lib.check(dev_call(), name)
                    k5.append(lib.dk_model_last_kernel_ms(h5.ptr))
                dstep(); k5.clear()
                ms5, _ = timed(dstep, 1)
                ms5h, _ = timed(lambda: _lib.check(host_call(), name), 1)
                ach = allsum(fl * N5 / (k5[-1] * 1e-3) * 1e-12) / world
                same = bool(np.array_equal(o5.cpu().numpy(), o5_h, equal_nan=True))
                # which.max / which.min over the whole sharded set: (value, global index) pairs merged like the screen
                ov = o5_h
                li = int(np.nanargmax(ov)) if name == "akg" else int(np.nanargmin(ov))
                gv, gi = parallel.global_argmax(float(ov[li]) if name == "akg" else -float(ov[li]), li, a_lo, device="cuda")
                res[name] = {"value": N5T / (ms5 * 1e-3), "unit": UNIT, "e2e": {"value": N5T / (ms5h * 1e-3), "unit": UNIT,
                             "h2d_bytes_per_step": int(N5 * 20 * 8), "d2h_bytes_per_step": int(N5 * 8), "host_memory": "pageable"},
                             "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                                          "flops_per_candidate": fl, "kernel_ms": k5[-1], "per_gpu": True},
                             "host_equals_device_path": same, "finite": bool(np.isfinite(ov).all()),
                             "best": {"index": int(gi), "value": gv if name == "akg" else -gv}}
            if rank == 0:
                # agreement with the oracle on a 256-candidate sample (AKG: values; argmax of the sample)
                sys.path.insert(0, os.path.join(ROOT, "tests"))
                import parity_cases as pc
                from conftest import to_oracle
                from oracle import diceoptim as do
                ko = to_oracle(km5)
                xs = np.ascontiguousarray(x5_host[:, :256].T)
                xs_f = np.asfortranarray(xs)
                ga = np.empty(256); gq = np.empty(256)
                _lib.check(lib.dk_akg(h5.ptr, vp(xs_f), 256, 0.1, 1, vp(ga)), "dk_akg")
                _lib.check(lib.dk_quantile(h5.ptr, vp(xs_f), 256, QN01, 1, vp(gq), None), "dk_quantile")
                t0 = time.perf_counter()
                predX = pc.oracle_predX(ko)
                t_pred = time.perf_counter() - t0
                t0 = time.perf_counter()
                ra = np.array([do.AKG(xs[i], ko, new_noise_var=0.1, predX=predX) for i in range(256)])
                t_akg = time.perf_counter() - t0
                rq = do.kriging_quantile(xs, ko, beta=0.1)
                scale = float(np.max(np.abs(km5.y)))
                res["oracle_sample"] = {
                    "candidates": 256, "akg_argmax_equal": bool(int(np.argmax(ga)) == int(np.argmax(ra))),
                    "quantile_argmin_equal": bool(int(np.argmin(gq)) == int(np.argmin(rq))),
                    "akg_max_abs_err_over_scale": float(np.max(np.abs(ga - ra)) / scale),
                    "quantile_max_rel_err": float(np.max(np.abs(gq - rq) / np.abs(rq))),
                    "cpu_baseline_akg": {"value": 256 / t_akg, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                         "sample": f"256 candidates, oracle AKG with the design prediction hoisted "
                                                   f"({t_pred:.1f} s once per model; the reference repeats it per call)"}}
            extra["cfg5"][kind] = res
            km5.release()
        del x5, o5

        # ---- optional: the round-1 weak-scaling figure (10^7 candidates per GPU) ----
        if args.weak:
            xw = torch.from_numpy(candidate_rows(CFG, 0, NT, d, seed_extra=1 + rank)).to("cuda")
            stepw = lambda: _lib.check(lib.dk_screen_ei(h.ptr, vp(xw), 1, NT, plugin, 1, 1, 0, 1, vp(idx), vp(val), sptr), "screen")
            stepw()
            msw, _ = timed(stepw, args.steps)
            extra["weak"] = {"value": world * NT * args.steps / (msw * 1e-3), "unit": UNIT, "candidates_per_gpu_per_step": NT}
            del xw

        # ---- one process, all N GPUs: dk_multi_screen_ei with the whole host matrix (rank 0; the others wait on a
        #      HOST barrier: an NCCL barrier would keep a spinning kernel on the very GPUs rank 0 is about to use) ----
        barrier()
        if rank == 0:
            ndev = min(world, torch.cuda.device_count())
            from diceoptim_b200.km import MultiDeviceModel
            mdm = MultiDeviceModel(km, ndev)
            mm = mdm.ptr
            x_all = x_page if world == 1 else candidate_rows(CFG, 0, NT, d)
            im = np.empty(1, np.int64); vm = np.empty(1)
            call = lambda: _lib.check(lib.dk_multi_screen_ei(mm, vp(x_all), NT, plugin, 1, 1, 0, 1, vp(im), vp(vm)), "dk_multi_screen_ei")
            call()
            t0 = time.perf_counter(); call(); dtm = time.perf_counter() - t0
            extra["single_process"] = {"value": NT / dtm, "unit": UNIT, "devices": ndev, "timing": "host wall clock around one call (pageable host matrix, copies included)",
                                       "argmax_equals_sharded": bool(int(im[0]) == res_dev[0] and float(vm[0]) == res_dev[1])}
            mdm.close()
        if host_group is not None:
            dist.barrier(group=host_group)

    if rank == 0:
        # CPU baselines: the oracle port on bounded samples of the same workloads, host cores
        sample = args.cpu_sample
        xs = np.ascontiguousarray(x_page[:, :sample].T)
        dt, imax = cpu_oracle_rate(km, xs)
        line["cpu_baseline"] = {"value": sample / dt, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                "sample": f"first {sample} candidates, oracle EI + which.max "
                                          f"(numpy + scipy OpenBLAS dtrsm, chunks over all host threads), {dt:.1f} s"}
        idx_s = np.empty(1, dtype=np.int64); val_s = np.empty(1)
        xs_f = np.asfortranarray(xs)
        _lib.check(lib.dk_screen_ei(h.ptr, vp(xs_f), 0, sample, plugin, 1, 1, 0, 1, vp(idx_s), vp(val_s), None), "dk_screen_ei")
        line["cpu_baseline"]["argmax_matches_gpu"] = bool(int(idx_s[0]) == imax)
        if not args.no_extras:
            gs = 2048
            dtg = cpu_ei_grad_rate(km, xs[:gs])
            extra["ei_grad"]["cpu_baseline"] = {"value": gs / dtg, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                                "sample": f"{gs} candidates, oracle EI.grad one point per call (R/EI.grad.R:97-186), threads over points, {dtg:.1f} s"}
            qs = 256
            xbq = np.ascontiguousarray(xq_host[:, :qs * q].T).reshape(qs, q, km3.d)
            dtq, qref = cpu_qei_rate(CFG3, xbq, pl3)
            qgpu = oq_h[:qs]
            extra["qei"]["cpu_baseline"] = {"value": qs / dtq, "unit": "evals/s", "cores": os.cpu_count(), "kind": "port",
                                            "sample": f"{qs} batches, oracle qEI(fastCompute=FALSE) (R/qEI.R:119-194), one process per core, {dtq:.1f} s",
                                            "max_rel_err_gpu_vs_oracle": float(np.max(np.abs(qgpu - np.array(qref)) / np.maximum(np.abs(qref), 1e-12)))}
        if not args.no_extras:
            # latencies of the optimiser-sized shapes (rank 0, host calls, copies included; best of 5)
            def best_ms(fn, reps=5):
                b = 1e9
                for _ in range(reps):
                    t0 = time.perf_counter(); fn(); b = min(b, time.perf_counter() - t0)
                return 1e3 * b
            x16 = np.asfortranarray(xs[:16]); e16 = np.empty(16); g16 = np.empty((16, d), order="F")
            from diceoptim_b200.optim import device_polish
            x10 = xs[100:110].copy()
            lat = {"ei_16_candidates_ms": best_ms(lambda: _lib.check(lib.dk_ei(h.ptr, vp(x16), 16, plugin, 1, 1, 0, vp(e16), None, None), "ei")),
                   "ei_grad_16_candidates_ms": best_ms(lambda: _lib.check(lib.dk_ei_grad(h.ptr, vp(x16), 16, plugin, 1, 1, 0, vp(g16), vp(e16)), "grad")),
                   "polish_10_starts_maxit100_ms": best_ms(lambda: device_polish(km, "EI", x10, np.zeros(d), np.ones(d), plugin=plugin, maxit=100, device=local), 2)}
            hc = h.clone()
            xn = np.ascontiguousarray(np.random.default_rng(5).random((4, d)))
            t0 = time.perf_counter()
            for i in range(4):
                _lib.check(lib.dk_model_append(hc.ptr, vp(xn[i]), plugin, 0.0), "append")
            lat["model_append_ms_per_point"] = 1e3 * (time.perf_counter() - t0) / 4
            hc.close()
            extra["latency"] = lat
            # ---- the small BASELINE configs at their own sizes (cfg2: EI over 10^6 points; cfg1: EI on the 100 x 100 grid) ----
            small = {}
            for key, label in (("cfg2_hartman4_n200", "cfg2"), ("cfg1_branin2_n20", "cfg1")):
                kms = synth.config_model(key)
                hs = kms.handle(local)
                lib.dk_model_set_timing(hs.ptr, 1)
                xh = np.asfortranarray(synth.config_candidates(key))
                Ns = xh.shape[0]
                xd = torch.from_numpy(np.ascontiguousarray(xh.T)).to("cuda")          # d x N row-major = N x d column-major
                eo = torch.empty(Ns, dtype=torch.float64, device="cuda"); eh = np.empty(Ns)
                pls = float(np.min(kms.y)); ks = []
                for _ in range(6):
                    _lib.check(lib.dk_ei_dev(hs.ptr, vp(xd), Ns, pls, 1, 1, 0, vp(eo), None, None, sptr), "dk_ei_dev")
                    torch.cuda.synchronize(); ks.append(lib.dk_model_last_kernel_ms(hs.ptr))
                kt = min(ks[1:])
                ht = best_ms(lambda: _lib.check(lib.dk_ei(hs.ptr, vp(xh), Ns, pls, 1, 1, 0, vp(eh), None, None), "dk_ei"))
                assert np.array_equal(eo.cpu().numpy(), eh, equal_nan=True)
                fls = flops_per_candidate(kms.n, kms.p)
                ach = fls * Ns / (kt * 1e-3) * 1e-12
                small[label] = {"workload": key, "n": kms.n, "d": kms.d, "candidates": int(Ns), "value": Ns / (kt * 1e-3), "unit": UNIT,
                                "kernel": "k_predict_small" if lib.dk_model_small_kernel(hs.ptr) else "k_predict_fused_tma",
                                "kernel_ms": kt, "roofline": {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                                                              "flops_per_candidate": fls},
                                "e2e": {"value": Ns / (ht * 1e-3), "unit": UNIT, "ms": ht, "h2d_bytes": int(Ns * kms.d * 8),
                                        "d2h_bytes": int(Ns * 8), "host_memory": "pageable"}}
                del xd, eo
                kms.release()
            extra["small_designs"] = small
        if extra:
            line["extra"] = extra
        print(json.dumps(line), flush=True)
    if not args.no_extras:
        km3.release()
    km.release()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
