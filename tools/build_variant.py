"""Build an experimental variant of the library: python tools/build_variant.py <name> [-DFLAG ...]
-> diceoptim_b200/libdk_b200_x<name>.so (git-ignored).  Only dk_api.cu (forward / gradient kernels) is recompiled with the
flags; the other objects come from the regular build.  Time it with tools/ab_lib.py."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g
g.build()
name, flags = sys.argv[1], sys.argv[2:]
obj = os.path.join(g.OBJDIR, f"dk_api_x{name}.o")
nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
subprocess.run([nvcc] + g.NVCC_FLAGS + flags + ["-c", os.path.join(g.CSRC, "dk_api.cu"), "-o", obj], check=True)
others = [os.path.join(g.OBJDIR, f) for f in sorted(os.listdir(g.OBJDIR)) if f.endswith(".o") and not f.startswith("dk_api_x") and f != "dk_api.o"]
out = os.path.join(ROOT, "diceoptim_b200", f"libdk_b200_x{name}.so")
subprocess.run([nvcc, "-shared", "-o", out, obj] + others + ["-lgomp"], check=True)
print(out)
