"""CPU: the C-ABI library loads and exports every symbol include/dk_b200.h declares (no compute)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "dk_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dk_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    from diceoptim_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"{n} declared in dk_b200.h but not exported"
    # and the binding table covers the header
    assert set(names) <= set(_lib.SIGNATURES)


def test_version_and_argument_errors_without_gpu():
    from diceoptim_b200 import _lib
    lib = _lib.load()
    assert b"sm_100a" in lib.dk_version()
    assert lib.dk_model_create(None, None) != 0
    assert b"NULL" in lib.dk_last_error()
    assert lib.dk_ei(None, None, 0, 0.0, 1, 1, 0, None, None, None) != 0


def test_r_call_shim_compiles_against_declarations():
    """The .Call shim (r-package/src) type-checks against declarations of the R C API it uses
    (tests/r_stub: declarations only, no R runtime here) and names only symbols the header declares."""
    import subprocess
    srcs = [os.path.join(ROOT, "r-package", "src", f) for f in ("dk_rcall.c", "packagename_init.c")]
    cmd = ["gcc", "-fsyntax-only", "-Wall", "-Werror", "-DDK_HAVE_R", "-I" + os.path.join(ROOT, "tests", "r_stub"),
           "-I" + os.path.join(ROOT, "include")] + srcs
    subprocess.run(cmd, check=True)
    used = set(re.findall(r"\b(dk_[a-z_0-9]+)\s*\(", open(srcs[0]).read()))
    assert used <= set(_declared()) | {"dk_handle"}


def test_integration_doc_lists_every_registered_call_entry():
    """INTEGRATION.md section 1 is the registration file verbatim: every .Call entry registered in packagename_init.c is
    named there and defined in dk_rcall.c (round 1 shipped a 5-entry table next to an 11-entry file)."""
    init = open(os.path.join(ROOT, "r-package", "src", "packagename_init.c")).read()
    entries = set(re.findall(r'\{"(dkR_[a-z_]+)"', init))
    assert len(entries) >= 17
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    shim = open(os.path.join(ROOT, "r-package", "src", "dk_rcall.c")).read()
    rsrc = open(os.path.join(ROOT, "r-package", "R", "gpu_engine.R")).read()
    for e in entries:
        assert f'"{e}"' in doc, f"{e} registered but missing from INTEGRATION.md"
        assert re.search(r"SEXP\s+" + e + r"\s*\(", shim), f"{e} registered but not defined in dk_rcall.c"
    for e in set(re.findall(r'\.Call\("(dkR_[a-z_]+)"', rsrc)):
        assert e in entries, f"{e} called from R but not registered"
