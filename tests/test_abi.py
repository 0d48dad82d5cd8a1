"""The C-ABI library loads and exports every symbol include/*.h declares (no compute calls: runs without a GPU)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(orb[xm]_[a-z0-9_]+)\s*\(", src)))


def test_exports():
    from orb_slam_b200 import build
    L = ctypes.CDLL(build.build())
    names = _declared("orbx.h") + _declared("orbm.h")
    assert len(names) > 35
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_no_device_is_an_error_not_a_fallback():
    """Without a usable device orbx_create / orbm_create must fail loudly (no CPU path exists in the product)."""
    import torch
    if torch.cuda.is_available():
        return
    from orb_slam_b200 import Matcher, ORBextractor, OrbmError, OrbxError
    for ctor, exc in ((ORBextractor, OrbxError), (Matcher, OrbmError)):
        try:
            ctor()
        except exc as e:
            assert e.code == -2
        else:
            raise AssertionError("creation succeeded without a GPU")


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "orb_slam_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".h")):
                s = open(os.path.join(dp, f), errors="ignore").read()
                assert "import oracle" not in s and "from oracle" not in s and "liborb_oracle" not in s, f


def test_headers_are_plain_c(tmp_path):
    """include/orbx.h and include/orbm.h compile as C99 (no C++ in the ABI headers) and every declared function links."""
    import subprocess
    from orb_slam_b200 import build
    lib = build.build()
    src = tmp_path / "t.c"
    names = _declared("orbx.h") + _declared("orbm.h")
    body = "\n".join(f"  p[{i}] = (void*){n};" for i, n in enumerate(names))
    src.write_text('#include "orbx.h"\n#include "orbm.h"\nint main(void) {\n  void* p[%d];\n%s\n  return p[0] == 0;\n}\n'
                   % (len(names), body))
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-Wno-pedantic", "-I", os.path.join(ROOT, "include"), str(src),
                    "-o", str(tmp_path / "t"), "-L", os.path.dirname(lib), "-lorbx",
                    f"-Wl,-rpath,{os.path.dirname(lib)}"], check=True)
