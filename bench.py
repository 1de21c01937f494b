#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 acquisition engine (contract in the task brief).

Workload (BASELINE.json configs[3], the one the metric is quoted on): max_EI candidate screening,
km matern5_2, n = 2000, d = 10, N = 10^7 uniform candidates per GPU per step (weak scaling: every
rank screens its own 10^7 candidates, the model is replicated, the only collective is the
all-gather of each rank's (value, index) pair for the global argmax).

  value : candidates/s, candidates already resident in HBM when the timed region starts
          (dk_screen_ei with a device pointer: fused EI kernel + device which.max + 16 B result).
  e2e   : the same call with HOST buffers (pinned): H2D of the 800 MB candidate matrix and D2H of
          the result inside the timed region — what R's .Call shim would see.
  roofline : FP64 tensor pipe; achieved = algorithmic flops (n^2 + 4n + 2np per candidate,
          SURVEY 8(d)) / CUDA-event time of the fused solve kernel, peak = cuBLAS DGEMM measured
          on this pool (profiles/dgemm_peak_r01.json; MEASURED_PEAKS.json has no FP64 entry).
  cpu_baseline : the oracle port (numpy + scipy/OpenBLAS dtrsm) on a bounded candidate sample.

`--impl reference` times the reference's CPU algorithm (the oracle port: R is not installed, the
arithmetic lives in DiceKriging) on the host cores for the same metric and config.
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
METRIC = "EI max_EI-screen candidates/sec at n=2000,d=10 (km matern5_2)"
UNIT = "candidates/s"


def flops_per_candidate(n, p):
    return n * n + 4 * n + 2 * n * p


def fp64_peak():
    """Measured cuBLAS DGEMM TFLOP/s on this pool's B200 (tools/microbench/dgemm_peak.py)."""
    path = os.path.join(ROOT, "profiles", "dgemm_peak_r01.json")
    try:
        return float(json.load(open(path))["burst_tflops"]), "measured cuBLAS DGEMM 8192^3 (profiles/dgemm_peak_r01.json)"
    except Exception:
        return 37.0, "fallback: DMMA issue peak 64 FMA/clk/SM x 148 SM x 1.965 GHz"


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


def run_reference(args):
    """The reference's CPU algorithm for the path (oracle port) on all host threads; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    from diceoptim_b200 import synth
    km = synth.config_model(CFG)
    sample = args.cpu_sample
    x = synth.config_candidates(CFG, N=sample)
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
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": CFG, "n": km.n, "d": km.d, "kernel": "matern5_2", "candidates_per_step": sample,
                   "note": "bounded sample of the 10^7-candidate workload; the rate is per candidate"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{sample} candidates/step x {args.steps} steps, numpy + scipy OpenBLAS dtrsm, chunks over {cores} threads"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--candidates", type=int, default=10_000_000, help="candidates per GPU per step")
    ap.add_argument("--cpu-sample", type=int, default=64000, help="candidates in the CPU-baseline sample")
    ap.add_argument("--no-extras", action="store_true", help="skip the EI+grad and vector-return side measurements")
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
    from diceoptim_b200 import synth, _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _lib.load()

    km = synth.config_model(CFG)
    n, d, p = km.n, km.d, km.p
    h = km.handle(local)
    lib.dk_model_set_timing(h.ptr, 1)
    N = args.candidates
    plugin = float(np.min(km.y))

    # candidates: this rank's shard, generated once; pinned host copy for e2e, device copy for value
    x_host = torch.empty((d, N), dtype=torch.float64, pin_memory=True)   # column-major N x d == row-major d x N
    rng = np.random.default_rng([synth.CONFIGS[CFG]["cand_seed"], rank])
    xv = x_host.numpy()
    for k in range(d):
        xv[k] = rng.random(N)
    x_dev = x_host.to("cuda", non_blocking=False)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sptr = C.c_void_p(stream.cuda_stream)
    idx = np.empty(1, dtype=np.int64); val = np.empty(1)
    from diceoptim_b200 import parallel

    def global_argmax():
        # all-gather of one (value, global index) pair per rank; first maximum wins
        v, gi = parallel.global_argmax(float(val[0]), int(idx[0]), rank * N, device="cuda")
        return gi, v

    def step(on_device):
        ptr = C.c_void_p(x_dev.data_ptr()) if on_device else C.c_void_p(x_host.data_ptr())
        _lib.check(lib.dk_screen_ei(h.ptr, ptr, 1 if on_device else 0, N, plugin, 1, 1, 0, 1,
                                    C.c_void_p(idx.ctypes.data), C.c_void_p(val.ctypes.data), sptr), "dk_screen_ei")
        return global_argmax()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(on_device, steps, sample_clocks=False):
        sampler = ClockSampler(local) if sample_clocks else None
        kernel_ms = []
        barrier()
        if sampler:
            sampler.start()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        l0 = h.launches()
        e0.record(stream)
        for _ in range(steps):
            res = step(on_device)
            kernel_ms.append(lib.dk_model_last_kernel_ms(h.ptr))
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, kernel_ms, h.launches() - l0, res, clocks

    for _ in range(args.warmup):
        step(True)
    ms_dev, kernel_ms, launches, res_dev, clocks = timed(True, args.steps, sample_clocks=True)
    step(False)
    ms_e2e, _, _, res_e2e, _ = timed(False, args.steps)
    value = world * N * args.steps / (ms_dev * 1e-3)
    e2e = world * N * args.steps / (ms_e2e * 1e-3)
    assert res_dev == res_e2e, (res_dev, res_e2e)

    peak, peak_src = fp64_peak()
    k_ms = sum(kernel_ms) / len(kernel_ms)
    achieved = flops_per_candidate(n, p) * N / (k_ms * 1e-3) * 1e-12
    traffic = None
    try:
        # ncu capture (profiles/ncu_k_predict_fused_r01.md) gives DRAM bytes per candidate; one launch screens N of them
        traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))["k_predict_fused_bytes_per_candidate"] * N
    except Exception:
        pass

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": CFG, "n": n, "d": d, "p": p, "kernel": "matern5_2", "type": "UK",
                   "candidates_per_gpu_per_step": N, "parallelism": f"candidates sharded x{world}, model replicated",
                   "l2": "inputs (800 MB/step) and per-CTA panels (296 MB) exceed the 126 MB L2; no flush needed",
                   "argmax": {"index": res_dev[0], "value": res_dev[1]}},
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(N * d * 8), "d2h_bytes_per_step": 16,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                     "traffic": traffic, "kernel": "k_predict_fused", "kernel_ms": k_ms,
                     "flops_per_candidate": flops_per_candidate(n, p), "peak_source": peak_src},
        "clocks": clocks,
    }

    if rank == 0 and not args.no_extras:
        # side measurements on a smaller batch: EI + gradient (adjoint form) and full-vector return
        Ng = 1 << 20
        xg = x_dev[:, :Ng].contiguous()
        grad = torch.empty((d, Ng), dtype=torch.float64, device="cuda")
        eiv = torch.empty(Ng, dtype=torch.float64, device="cuda")
        def gstep():
            _lib.check(lib.dk_ei_grad_dev(h.ptr, C.c_void_p(xg.data_ptr()), Ng, plugin, 1, 1, 0,
                                          C.c_void_p(grad.data_ptr()), C.c_void_p(eiv.data_ptr()), sptr), "dk_ei_grad_dev")
        gstep(); torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream); gstep(); gstep(); e1.record(stream); torch.cuda.synchronize()
        gms = e0.elapsed_time(e1) / 2
        gfl = (2 * n * n + 2 * n * d + 4 * n) * Ng / (gms * 1e-3) * 1e-12
        line["extra"] = {"ei_grad_candidates_per_s": Ng / (gms * 1e-3), "ei_grad_batch": Ng,
                         "ei_grad_tflops_algorithmic": gfl, "ei_grad_frac_of_peak": gfl / peak}
        # qEI (q = 4) on BASELINE config 3: rosenbrock4 km, n = 500, 10^5 batches; device-resident batches
        km3 = synth.config_model("cfg3_rosenbrock4_n500")
        h3 = km3.handle(local)
        Bq, q = 100_000, 4
        xq = torch.rand((km3.d, Bq * q), dtype=torch.float64, device="cuda")
        oq = torch.empty(Bq, dtype=torch.float64, device="cuda")
        pl3 = float(np.median(km3.y))
        def qstep():
            _lib.check(lib.dk_qei_dev(h3.ptr, C.c_void_p(xq.data_ptr()), Bq, q, pl3, 1, 1, C.c_void_p(oq.data_ptr()), sptr),
                       "dk_qei_dev")
        qstep(); torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream); qstep(); qstep(); qstep(); e1.record(stream); torch.cuda.synchronize()
        qms = e0.elapsed_time(e1) / 3
        line["extra"].update({"qei_q4_evals_per_s": Bq / (qms * 1e-3), "qei_batches": Bq, "qei_model": "cfg3 n=500 d=4 matern5_2",
                              "qei_ms": qms, "qei_nan": int(torch.isnan(oq).sum().item())})
        km3.release()
        # AKG + kriging.quantile on BASELINE config 5: d = 20, n = 5000, noise.var = 0.1
        km5 = synth.config_model("cfg5_noisy20_n5000")
        h5 = km5.handle(local)
        lib.dk_model_set_timing(h5.ptr, 1)
        Na = 1 << 16
        xa = torch.rand((km5.d, Na), dtype=torch.float64, device="cuda")
        oa = torch.empty(Na, dtype=torch.float64, device="cuda")
        for name, call in (("akg", lambda: lib.dk_akg_dev(h5.ptr, C.c_void_p(xa.data_ptr()), Na, 0.1, 1, C.c_void_p(oa.data_ptr()), sptr)),
                           ("quantile", lambda: lib.dk_quantile_dev(h5.ptr, C.c_void_p(xa.data_ptr()), Na, -1.2815515655446004, 1,
                                                                    C.c_void_p(oa.data_ptr()), None, sptr))):
            _lib.check(call(), name); torch.cuda.synchronize()
            _lib.check(call(), name); torch.cuda.synchronize()
            ms5 = lib.dk_model_last_kernel_ms(h5.ptr)
            fl5 = (2 if name == "akg" else 1) * km5.n ** 2 * Na / (ms5 * 1e-3) * 1e-12
            line["extra"].update({f"{name}_n5000_candidates_per_s": Na / (ms5 * 1e-3), f"{name}_n5000_tflops_algorithmic": fl5,
                                  f"{name}_n5000_finite": bool(torch.isfinite(oa).all().item())})
        km5.release()

    if rank == 0:
        # CPU baseline: the oracle port on a bounded sample of the same workload, host cores
        sample = args.cpu_sample
        xs = np.ascontiguousarray(x_host.numpy()[:, :sample].T)
        dt, imax = cpu_oracle_rate(km, xs)
        line["cpu_baseline"] = {"value": sample / dt, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                "sample": f"first {sample} candidates of rank 0's shard, oracle EI + which.max "
                                          f"(numpy + scipy OpenBLAS dtrsm, chunks over all host threads), {dt:.1f} s"}
        # cheap end-to-end parity check of the screen on that sample
        idx_s = np.empty(1, dtype=np.int64); val_s = np.empty(1)
        xs_f = np.asfortranarray(xs)
        _lib.check(lib.dk_screen_ei(h.ptr, C.c_void_p(xs_f.ctypes.data), 0, sample, plugin, 1, 1, 0, 1,
                                    C.c_void_p(idx_s.ctypes.data), C.c_void_p(val_s.ctypes.data), None), "dk_screen_ei")
        line["cpu_baseline"]["argmax_matches_gpu"] = bool(int(idx_s[0]) == imax)
        print(json.dumps(line), flush=True)
    km.release()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
