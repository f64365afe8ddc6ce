"""The C-ABI library loads and exports every symbol include/jtsm_b200.h declares (no GPU)."""
import ctypes

from jtsm_b200 import _lib


def test_header_parses():
    fns = _lib.declared_functions()
    assert "jtsm_roi_pool_forward" in fns
    assert len(fns) >= 8
    for name, (ret, args) in fns.items():
        assert name.startswith("jtsm_")


def test_library_exports_every_declared_symbol(lib_built):
    h = ctypes.CDLL(lib_built)
    for name in _lib.declared_functions():
        assert hasattr(h, name), "libjtsm_b200.so does not export %s" % name


def test_no_undeclared_exports(lib_built):
    import subprocess

    out = subprocess.run(["nm", "-D", "--defined-only", lib_built], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    exported = {s for s in exported if s.startswith("jtsm_")}
    assert exported == set(_lib.declared_functions())


def test_host_only_entry_points(lib_built):
    L = _lib.lib()
    assert L.jtsm_abi_version() == 1
    assert L.jtsm_error_string(0) == b"ok"
    assert L.jtsm_error_string(3) == b"workspace too small"
    assert L.jtsm_roi_pool_workspace_bytes(1, 2000, 7, 7) > 2000 * 7 * 4 * 2
    v = ctypes.c_int(-1)
    assert L.jtsm_get_tuning(b"roi_pool_splits", ctypes.byref(v)) == 0 and v.value == 0
    assert L.jtsm_set_tuning(b"no_such_key", 1) == 1


def test_ops_refuse_cpu_tensors(lib_built):
    import pytest
    import torch

    from jtsm_b200 import ops

    with pytest.raises(RuntimeError, match="CUDA-only"):
        ops.roi_pool_forward(torch.zeros(1, 1, 4, 4), torch.zeros(1, 5), 1.0, (2, 2))
