"""The C-ABI library loads and exports every symbol include/ramp_b200.h declares; without a
GPU it fails loudly instead of falling back to a CPU path.  No compute calls here."""

import ctypes as C
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "ramp_b200.h"


def _declared():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(ramp_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_are_exported(built):
    from ramp_b200 import engine

    lib = engine.load_library()
    names = _declared()
    assert len(names) >= 15 and "ramp_check_batch" in names and "ramp_check_edges" in names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", str(engine.LIB_PATH)], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (ramp_[a-z_0-9]+)", out))
    assert set(names) <= exported
    assert exported <= set(names), f"exported but undeclared: {exported - set(names)}"
    assert lib.ramp_abi_version() == 1


def test_struct_sizes_match_the_header(built):
    """ctypes mirrors vs a C translation unit that prints sizeof() of the header's structs."""
    from ramp_b200 import desc as D

    src = ('#include <stdio.h>\n#include "ramp_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu", sizeof(RampJoint),'
           "sizeof(RampGeom), sizeof(RampQMap), sizeof(RampModelDesc), sizeof(RampSubspace), sizeof(RampSpaceDesc));return 0;}")
    exe = ROOT / "tests" / "_sizes.out"
    try:
        subprocess.run(["/usr/bin/gcc", "-x", "c", "-", "-I", str(ROOT / "include"), "-o", str(exe)], input=src, text=True, check=True)
        sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    finally:
        exe.unlink(missing_ok=True)
    assert sizes == [C.sizeof(t) for t in (D.RampJoint, D.RampGeom, D.RampQMap, D.RampModelDesc, D.RampSubspace, D.RampSpaceDesc)]


def test_no_cpu_fallback_without_a_device(built):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from ramp_b200 import engine
    from ramp_b200.desc import ModelDesc
    from ramp_b200.exceptions import RampGpuError
    from ramp_b200.robot_model import load_robot_model
    from ramp_b200.robot_state import RobotState

    rm = load_robot_model("zoo::rrr/configs")
    lib = engine.load_library()
    d = ModelDesc(rm.model, rm.collision_model)
    handle = C.c_void_p()
    rc = lib.ramp_model_create(C.byref(d.struct), 0, C.byref(handle))
    assert rc == -3 and not handle  # RAMP_ERR_CUDA
    assert b"no CPU fallback" in lib.ramp_last_error()
    with pytest.raises(RampGpuError):
        RobotState(rm).check_collision()


def test_product_never_imports_the_oracle():
    for f in (ROOT / "ramp_b200").rglob("*.py"):
        assert not re.search(r"^\s*(from|import)\s+oracle\b", f.read_text(), flags=re.M), f
    for f in (ROOT / "ramp_b200" / "csrc").iterdir():
        assert "oracle" not in f.read_text(), f
