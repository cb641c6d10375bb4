"""The C-ABI library loads and exports every symbol include/crm_b200.h declares; host-only entry points
(no device) answer; the simulation entry points fail loudly without a GPU (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from cr_mech_coli_b200 import _capi

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    text = (ROOT / "include" / "crm_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(crm_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported():
    lib = _capi.product()
    names = declared_symbols()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib.lib, n)]
    assert not missing, missing
    assert lib.fn("abi_version")() == 3


def test_struct_layouts_match_header():
    assert C.sizeof(_capi.CrmConfig) == 64
    assert _capi.ROD_PARAMS_DTYPE.itemsize == 100 and _capi.IDENT_DTYPE.itemsize == 24


def test_save_schedule_default():
    # default Configuration: 12 snapshots at round(n*100/11/0.1) (reference src/simulation.rs:239-249)
    lib = _capi.product()
    its, last = lib.save_schedule(_capi.make_config())
    assert last == 1000
    assert list(its) == [0, 91, 182, 273, 364, 455, 545, 636, 727, 818, 909, 1000]
    its, last = lib.save_schedule(_capi.make_config(n_saves=0, t_max=2.0))
    assert list(its) == [0, 20] and last == 20
    with pytest.raises(_capi.SimulationError, match="TimeError"):
        lib.save_schedule(_capi.make_config(dt=0.0))


def test_colors_through_abi():
    lib = _capi.product()
    for counter in (0, 1, 2, 250, 251, 63000, 251**3 - 1):
        rgb = (C.c_uint8 * 3)()
        lib.fn("counter_to_color")(counter, C.byref(rgb))
        assert lib.fn("color_to_counter")(C.byref(rgb)) == counter


def test_no_cpu_fallback_without_gpu():
    lib = _capi.product()
    if lib.fn("device_count")() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(_capi.SimulationError, match="no CUDA device"):
        _capi.Sim(lib, _capi.make_config())
    import cr_mech_coli_b200 as crm
    agents = crm.generate_agents(2, crm.AgentSettings(), crm.Configuration())
    with pytest.raises(crm.SimulationError):
        crm.run_simulation_with_agents(crm.Configuration(t_max=1.0), agents)
    # page-locked buffers come from the CUDA runtime: no device, no silent pageable stand-in
    with pytest.raises(MemoryError):
        lib.pinned_empty((4, 8, 3), "float32")


def test_product_package_does_not_import_oracle():
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|libcrm_oracle|CDLL\([^)]*oracle", re.M)
    for py in (ROOT / "cr_mech_coli_b200").glob("*.py"):
        text = "\n".join(l for l in py.read_text().splitlines() if not l.lstrip().startswith(("#", '"', "to drive", "is loaded")))
        assert not pat.search(text.replace("``oracle/libcrm_oracle.so``", "")), py
    for src in (ROOT / "cr_mech_coli_b200" / "csrc").glob("*.c*"):
        assert "oracle/" not in "\n".join(l for l in src.read_text().splitlines() if "//" not in l), src
