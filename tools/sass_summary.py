"""SASS evidence per hot kernel: counts of DMMA / UBLKCP (bulk async copy, TMA 1-D) / SYNCS (mbarrier) / LDS.128 / BAR and the
inner loop of the GEMM phase.  usage: python tools/sass_summary.py > profiles/sass_r02.md  (cuobjdump on the built .so; no GPU)"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "diceoptim_b200", "libdk_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)[1:]
HOT = ["k_predict_fused_tmaILi3ELb0", "k_predict_narrow", "k_grad_fusedILi3", "k_grad_narrow", "k_akg_fusedILb0", "k_qei_moments", "k_qei_cdfE",
       "k_qei_cdf_gen", "k_qei_moments_gen", "k_topk_pass", "k_gen_uniform", "k_polish_prepare", "k_predict_small", "k_qei_grad_moments_gen",
       "k_qei_grad_cdf_gen"]
pat = {"DMMA": r"\bDMMA", "UBLKCP": r"\bUBLKCP", "SYNCS": r"\bSYNCS", "LDS.128": r"\bLDS\.128", "BAR.SYNC": r"\bBAR\.SYNC", "DFMA": r"\bDFMA",
       "DADD/DMUL": r"\bD(ADD|MUL)", "MUFU": r"\bMUFU", "UTMALDG/UTC*MMA": r"\b(UTMALDG|UTC\w*MMA|LDTM|STTM)"}
print("# SASS evidence, round 2 (`cuobjdump -sass diceoptim_b200/libdk_b200.so`, sm_100a)\n")
print("`tcgen05.mma` has no f64 kind, so the FP64 contraction is `mma.sync.m8n8k4.f64` = SASS `DMMA.8x8x4`; the GEMM-phase tiles arrive by")
print("1-D bulk asynchronous copies (`cp.async.bulk` = SASS `UBLKCP`) completing on mbarriers (`SYNCS`), from this GPU's memory only, no multicast.\n")
print("| kernel | " + " | ".join(pat) + " | instructions |")
print("|---|" + "---|" * (len(pat) + 1))
loops = {}
seen = set()
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    if name in seen:
        continue
    seen.add(name)
    hit = [h for h in HOT if h in name]
    if not hit:
        continue
    body = f
    lines = [l for l in body.split("\n") if re.search(r"/\*[0-9a-f]{4,6}\*/", l)]
    row = [str(len(re.findall(p, body))) for p in pat.values()]
    short = re.sub(r"^_ZN2dk\d+|^_ZN\d+_GLOBAL__N__\w+?\d+(?=k_)", "", name)[:48]
    print(f"| `{short}` | " + " | ".join(row) + f" | {len(lines)} |")
    if hit[0] == "k_predict_fused_tmaILi3ELb0":
        # the densest run of DMMAs: one k4 step of mma_stage (16 DMMA fed by 4 LDS.128)
        idx = [i for i, l in enumerate(lines) if "DMMA" in l]
        best = max(range(len(idx) - 15), key=lambda j: -(idx[j + 15] - idx[j]))
        loops["k_predict_fused_tma<matern5_2>: one k4 step of the GEMM stage (16 DMMA.8x8x4 on a 32 x 32 warp tile)"] = lines[idx[best] - 6: idx[best + 15] + 2]
    if hit[0] == "k_predict_fused_tmaILi3ELb0":
        ub = [i for i, l in enumerate(lines) if "UBLKCP" in l]
        if ub:
            loops["k_predict_fused_tma<matern5_2>: stage issue (thread 0): arrive.expect_tx + two bulk copies (A tile 16 KB, B tile 8 KB)"] = lines[max(0, ub[0] - 8): ub[min(1, len(ub) - 1)] + 2]
for title, ls in loops.items():
    print(f"\n## {title}\n\n```")
    for l in ls:
        print(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip())
    print("```")
