"""`not gpu`: the C-ABI library loads, exports every symbol include/rtb200.h declares, agrees with the Python mirror on
struct sizes, and fails loudly (no CPU fallback) without a CUDA device.  No compute calls."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "rtb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", src)) - {"rt_update_cb"})


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.abi.load_library()
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rtb200.h but not exported by librtb200.so"
    assert sorted(pkg.abi.EXPORTS) == names, "abi.EXPORTS out of sync with include/rtb200.h"


def test_struct_sizes(pkg):
    lib = pkg.abi.load_library()
    assert lib.rt_struct_size(b"rt_triangle") == pkg.abi.triangle_dt.itemsize == 88
    assert lib.rt_struct_size(b"rt_sphere") == 24 and lib.rt_struct_size(b"rt_instance_xform") == 36
    assert lib.rt_struct_size(b"nope") == 0
    assert b"sm_100a" in lib.rt_version()


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(pkg.RtError) as e:
        pkg.Renderer(device=0)
    assert e.value.code == pkg.abi.RT_ERR_CUDA


def test_missing_extension_fails_loudly(pkg, tmp_path):
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.abi.load_library(str(tmp_path / "librtb200.so"))


def test_product_does_not_link_the_oracle(pkg):
    import subprocess
    out = subprocess.run(["ldd", pkg.abi.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "emul" not in out
    syms = subprocess.run(["nm", "-D", "--defined-only", pkg.abi.LIB_PATH], capture_output=True, text=True).stdout
    assert "port_render" not in syms and "ref_render" not in syms and "emul_render" not in syms
