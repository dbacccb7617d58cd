"""CPU: the C-ABI library loads without a GPU and exports every symbol include/vrgpu.h declares; calls fail loudly
(VRGPU_ENODEV) instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "vrgpu.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(vrgpu_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported():
    import vrgpu
    lib = vrgpu.load_library()
    names = _declared()
    assert "vrgpu_fof3d" in names and "vrgpu_search_full_set" in names
    for s in names:
        assert hasattr(lib, s), f"libvrgpu.so does not export {s}"
    assert set(vrgpu.EXPORTED_SYMBOLS) <= set(names)
    assert lib.vrgpu_version().decode().endswith("sm_100a")


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import vrgpu
    with pytest.raises(vrgpu.VrgpuError) as e:
        vrgpu.Context(0)
    assert e.value.code == 1  # VRGPU_ENODEV


def test_product_does_not_reference_oracle():
    """The product path must never import / link / call the checkers under oracle/."""
    pkg = os.path.join(ROOT, "velociraptor-stf_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cxx", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "vroracle" not in txt and "liboracle" not in txt and "libvrref" not in txt, f


def test_struct_layouts_match_header():
    """ctypes mirrors have the sizes the C compiler gives the header's structs."""
    import subprocess
    import tempfile
    import vrgpu
    src = '#include <stdio.h>\n#include "vrgpu.h"\nint main(){printf("%zu %zu %zu %zu\\n",sizeof(vrgpu_view),sizeof(vrgpu_unbind_opts),sizeof(vrgpu_search_opts),sizeof(vrgpu_stats));return 0;}\n'
    with tempfile.TemporaryDirectory() as td:
        c = os.path.join(td, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(td, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        sizes = [int(x) for x in subprocess.check_output([exe]).split()]
    assert sizes == [C.sizeof(vrgpu.View), C.sizeof(vrgpu.UnbindOpts), C.sizeof(vrgpu.SearchOpts), C.sizeof(vrgpu.Stats)]
