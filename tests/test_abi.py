"""The C-ABI library loads and exports every symbol include/chrono_b200.h declares (no compute calls: CPU-only)."""
import ctypes
import os
import re

from chrono_wisc_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "chrono_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cb200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_python_binding_covers_header():
    assert sorted(_lib._SIGNATURES) == declared_symbols()
    _lib.load()


def test_struct_sizes_match_header():
    # sizes computed from the C declarations: any drift between header and ctypes mirror shows up here
    assert ctypes.sizeof(_lib.cb200_settings) == 8 + 4 + 12 + 24 + 8 + 4 + 4 + 8 + 24 + 16 + 32 + 24 + 24 + 4 + 28
    assert ctypes.sizeof(_lib.cb200_counts) == 16 * 4 + 9 * 8 + 16


def test_default_settings_are_reference_defaults():
    lib = _lib.load()
    s = _lib.cb200_settings()
    lib.cb200_default_settings(ctypes.byref(s))
    # src/chrono_multicore/ChSettings.h:41-48,96-131
    assert list(s.bins_per_axis) == [10, 10, 10] and s.grid_density == 5 and s.narrowphase_algorithm == 2
    assert s.contact_force_model == 1 and s.tangential_displ_mode == 1 and s.use_material_properties == 1
    assert s.alpha == 0.2 and s.contact_recovery_speed == 0.6 and s.max_iteration_sliding == 100
    assert s.tolerance == 1e-4 and s.min_slip_vel == 1e-4


def test_no_cpu_fallback_without_device():
    lib = _lib.load()
    if lib.cb200_device_count() > 0:
        return
    h = ctypes.c_void_p()
    rc = lib.cb200_create(0, ctypes.byref(h))
    assert rc != 0 and b"no CUDA device" in lib.cb200_last_error()


def test_oracle_settings_mirror_matches_the_boundary_struct():
    """The checkers keep their own ctypes mirror of cb200_settings (they never load the product library); it must stay
    field-for-field identical to the product's."""
    from chrono_wisc_b200 import _lib
    from oracle import pyoracle

    a, b = _lib.cb200_settings, pyoracle.OrcSettings
    assert [(n, t._type_ if hasattr(t, "_length_") else t, getattr(t, "_length_", 1)) for n, t in a._fields_] == \
           [(n, t._type_ if hasattr(t, "_length_") else t, getattr(t, "_length_", 1)) for n, t in b._fields_]
    import ctypes as C
    assert C.sizeof(a) == C.sizeof(b)
