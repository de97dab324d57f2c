"""CPU: the C-ABI library loads and exports every symbol include/ucnerf.h declares; the ctypes mirror of the
POD structs has the C layout; the product never touches oracle/ and fails loudly without a GPU."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "ucnerf.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(ucn_[a-z0-9_]+)\s*\(", hdr)))


def test_header_symbols_exported():
    from ucnerf_b200 import build, lib as L
    build.build()
    lib = C.CDLL(L.LIB_PATH)
    declared = _declared()
    assert len(declared) >= 17
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/ucnerf.h but not exported"
    assert sorted(L.exported_symbols()) == declared, "ctypes signature table out of sync with the header"
    L.load()
    assert L.load().ucn_version() >= 100


def test_struct_layout_matches_c(tmp_path):
    """sizeof/offsetof of the ctypes mirrors == the C compiler's view of include/ucnerf.h."""
    from ucnerf_b200 import lib as L
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "ucnerf.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(ucn_frame),offsetof(ucn_frame,src_w2c),offsetof(ucn_frame,nearfar),offsetof(ucn_frame,feat),'
                   'sizeof(ucn_weights),sizeof(ucn_state_dict),offsetof(ucn_state_dict,film_w));return 0;}\n')
    exe = tmp_path / "layout"
    subprocess.run(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(L.Frame), L.Frame.src_w2c.offset, L.Frame.nearfar.offset, L.Frame.feat.offset,
            C.sizeof(L.Weights), C.sizeof(L.StateDict), L.StateDict.film_w.offset]
    assert got == want


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "uc-nerf_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f
                assert "libucnerf_oracle" not in txt and not re.search(r"#include\s+[<\"].*oracle", txt), f


def test_no_cpu_fallback():
    import torch
    from ucnerf_b200 import lib as L
    from ucnerf_b200.frame import PackedFrame
    from ucnerf_b200 import renderer
    with pytest.raises(L.UcnError):
        PackedFrame("cpu")
    with pytest.raises(L.UcnError):
        renderer.raw2outputs(torch.zeros(2, 3, 4), torch.zeros(2, 3), None)
    if not torch.cuda.is_available():
        assert L.load().ucn_check_device() != 0


def test_argument_validation_without_gpu():
    """Host-side validation runs before any launch, so it is testable on CPU."""
    from ucnerf_b200 import lib as L
    lib = L.load()
    fr = L.Frame()
    fr.S = 128                                   # not a multiple of 3 (reference crashes for it too, Q1)
    one = C.c_void_p(16)
    rc = lib.ucn_raygen_sample(C.byref(fr), None, 0, 4, one, None, 0, one, None, None, None, None)
    assert rc == -1 and b"multiple of 3" in lib.ucn_last_error()
    assert lib.ucn_weights_f32_bytes(3) >= 4 * 143000      # 177 034 params minus the 3 unused layers
    assert lib.ucn_weights_bf16_bytes(3) > 2 * 1136 * 128
