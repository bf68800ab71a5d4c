"""CPU test: libfxgpu.so loads without a GPU and exports every symbol include/fxgpu.h declares;
computing entry points fail loudly (no CPU fallback) when no CUDA device is usable."""
import ctypes as C
import os
import re

import helpers


def declared_symbols():
    text = open(os.path.join(helpers.ROOT, "include", "fxgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fxg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(built):
    from fxtract_b200 import lib
    L = lib.load_library()
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), "libfxgpu.so does not export %s" % n
    assert set(lib.EXPORTS) <= set(names)


def test_no_cpu_fallback_without_device(built):
    import torch
    if torch.cuda.is_available():
        return
    from fxtract_b200 import lib
    L = lib.load_library()
    cfg = lib.Config(0, 0, 0, 0, 0, 0, 0, None)
    h = C.c_void_p()
    rc = L.fxg_create(C.byref(cfg), C.byref(h))
    assert rc == -2 and not h.value            # FXG_ERR_CUDA
    assert b"no CPU fallback" in L.fxg_last_error(None)


def test_host_helpers_need_no_device(built):
    """Synthetic generator (host side) and the record cutter are pure host helpers."""
    import fxtract_b200 as fx
    from fxtract_b200 import lib
    spec = fx.SynthSpec(seed=7, first_record=5, n_records=100, read_len=50)
    data = lib.synth_host(spec)
    assert data.count(b"\n") == 400 and data.startswith(b"@SYN:")
    cut = lib.find_record_cut(data[:1000])
    assert 0 < cut <= 1000 and data[cut:cut + 1] == b"@" and data[cut - 1:cut] == b"\n"
    assert data[:cut].count(b"\n") % 4 == 0
