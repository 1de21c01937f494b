"""FP64 work of one qEI(q = 4) evaluation in k_qei_cdf, from ncu's executed-instruction counters, for bench.py's FP64-pipe
roofline entry of that kernel.  Run:
  ncu --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum \
      --clock-control none -k regex:k_qei_cdf --csv --log-file gpurun_out/qei_cdf_inst.csv python tools/qei_cdf_work.py
then  python tools/qei_cdf_work.py --parse gpurun_out/qei_cdf_inst.csv  writes profiles/qei_cdf_work.json."""
import csv, ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = 100_000

if len(sys.argv) > 2 and sys.argv[1] == "--parse":
    tot = {"dfma": 0.0, "dmul": 0.0, "dadd": 0.0}
    launches = set()
    for row in csv.reader(l for l in open(sys.argv[2]) if l.startswith('"')):
        if len(row) < 5 or "k_qei_cdf" not in " ".join(row):
            continue
        name = [c for c in row if c.startswith("smsp__sass_thread_inst_executed_op_")]
        if not name:
            continue
        for k in tot:
            if f"op_{k}_" in name[0]:
                tot[k] += float(row[-1].replace(",", ""))
        launches.add(row[0])
    nl = max(1, len(launches))
    flops = (2 * tot["dfma"] + tot["dmul"] + tot["dadd"]) / nl / B
    out = {"fp64_flops_per_eval_q4": flops, "fp64_pipe_peak_tflops": 34.0,
           "thread_inst_per_eval": {k: v / nl / B for k, v in tot.items()},
           "source": "ncu smsp__sass_thread_inst_executed_op_{dfma,dmul,dadd}_pred_on over k_qei_cdf at cfg3, 10^5 batches "
                     "(2 flops per DFMA); peak = measured DFMA issue peak, profiles/fp64_peaks_r01.md"}
    json.dump(out, open(os.path.join(ROOT, "profiles", "qei_cdf_work.json"), "w"), indent=1)
    print(json.dumps(out))
    sys.exit(0)

import numpy as np, torch
import __graft_entry__ as g
g.build()
from diceoptim_b200 import _lib, synth
lib = _lib.load()
km = synth.config_model("cfg3_rosenbrock4_n500"); h = km.handle(0)
x = torch.rand((km.d, B * 4), dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
o = torch.empty(B, dtype=torch.float64, device="cuda")
_lib.check(lib.dk_qei_dev(h.ptr, C.c_void_p(x.data_ptr()), B, 4, float(np.median(km.y)), 1, 1, C.c_void_p(o.data_ptr()), None), "qei")
torch.cuda.synchronize()
print("done")
