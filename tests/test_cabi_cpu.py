"""CPU checks of the drop-in boundary: the C-ABI library builds, loads and exports every symbol
include/mjsim_b200.h declares; the ctypes mirrors match the C structs; no CPU fallback exists."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def cabi():
    from mujoco_sim_b200 import lib

    if not lib.LIB_PATH.exists():
        lib.build()
    return lib.load()


def _declared_symbols():
    text = (ROOT / "include" / "mjsim_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mjs_[a-z_0-9]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(cabi):
    names = _declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(cabi, n), f"{n} declared in include/mjsim_b200.h but not exported"


def test_python_binding_covers_the_header():
    from mujoco_sim_b200 import lib

    assert sorted(lib.exported_symbols()) == _declared_symbols()


def test_struct_mirrors_match_the_oracle_build(oracle_lib):
    from mujoco_sim_b200.model.cstruct import MjsEnvState, MjsModel, MjsStepOut, MjsTask

    assert oracle_lib.oracle_sizeof_model() == ctypes.sizeof(MjsModel)
    assert oracle_lib.oracle_sizeof_task() == ctypes.sizeof(MjsTask)
    assert oracle_lib.oracle_sizeof_state() == ctypes.sizeof(MjsEnvState)
    assert oracle_lib.oracle_sizeof_out() == ctypes.sizeof(MjsStepOut)


def test_no_cpu_fallback(cabi):
    """Without a CUDA device mjs_create must fail loudly (-2) instead of falling back."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mujoco_sim_b200.model.tasks import build_pointmass_reach

    b = build_pointmass_reach()
    h = ctypes.c_void_p()
    rc = cabi.mjs_create(ctypes.byref(b.model), ctypes.byref(b.task), 4, 0, 0, 0, 1, ctypes.byref(h))
    assert rc != 0 and not h.value
    assert b"no CPU fallback" in cabi.mjs_last_error() or b"CUDA" in cabi.mjs_last_error()
    import mujoco_sim_b200 as ms

    with pytest.raises(ms.lib.MjsimError):
        ms.make("point_mass_reach", 4)


def test_product_never_imports_the_oracle():
    # no import / include of oracle/ anywhere in the package
    for p in (ROOT / "mujoco_sim_b200").rglob("*.py"):
        t = p.read_text()
        assert "from oracle" not in t and "import oracle" not in t, p
    for p in (ROOT / "mujoco_sim_b200" / "csrc").glob("*.cu*"):
        assert not re.search(r'#include\s+"[^"]*oracle', p.read_text()), p
