#!/usr/bin/env python
"""Achieved-error table per BASELINE config: GPU vs oracle, and both vs the long-double restatement.
Run on the GPU box:  python tools/parity_table.py > gpurun_out/parity_r02.md   (then copied to profiles/)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import __graft_entry__ as g
    g.build()
    import parity_cases as pc
    rows, notes = [], {}
    t0 = time.time()
    for name, fn in [
        ("cfg1", lambda: pc.ei_family("cfg1_branin2_n20", N=None, n_hi=64, n_grad_hi=64, full_grid=True)),
        ("cfg2", lambda: pc.ei_family("cfg2_hartman4_n200", N=8192, n_hi=64, n_grad_hi=32)),   # > 4736: the single-block kernel
        ("cfg3p", lambda: pc.ei_family("cfg3_rosenbrock4_n500", N=2048, n_hi=48, n_grad_hi=16)),
        ("cfg3", lambda: pc.qei_family(B=256)),
        ("cfg3_q5", lambda: pc.qei_family(B=32, q=5)),            # the general-q kernels (fixed-lattice CDFs on both sides)
        ("cfg3_q7", lambda: pc.qei_family(B=16, q=7)),
        ("cfg4", lambda: pc.ei_family("cfg4_sphere10_n2000", N=2048, n_hi=32, n_grad_hi=8)),
        ("cfg5_noise", lambda: pc.akg_family(noise_kind="noise_var", N=256, n_hi=32)),
        ("cfg5_nugget", lambda: pc.akg_family(noise_kind="nugget", N=256, n_hi=32)),
        ("akg_hi", lambda: pc.akg_hiprec()),
    ]:
        t = time.time()
        r, info = fn()
        rows += r
        notes[name] = {k: v for k, v in info.items() if not hasattr(v, "shape")}
        print(f"[{name}] {time.time() - t:.1f} s", file=sys.stderr, flush=True)
    f = lambda v: "—" if v is None else f"{v:.2e}"
    print("# Parity, round 2: achieved errors per BASELINE config\n")
    print("`relerr = max |a - ref| / max(|ref|, 1e-12 max|ref|)`; gradients per candidate relative to that candidate's largest")
    print("component.  *hiprec* = `oracle/hiprec.py` (long double + mpmath on the same float64 model slots) on a subset of the points.\n")
    print("| config | quantity | points | GPU vs oracle | GPU vs hiprec | oracle vs hiprec |")
    print("|---|---|---|---|---|---|")
    for r in rows:
        print(f"| {r['config']} | {r['quantity']} | {r['points']} | {f(r['gpu_vs_oracle'])} | {f(r['gpu_vs_hiprec'])} | {f(r['oracle_vs_hiprec'])} |")
    print(f"\nnotes: `{json.dumps(notes)}`; wall time {time.time() - t0:.0f} s")
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "parity_r02.json"), "w"), indent=1)


if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    main()
