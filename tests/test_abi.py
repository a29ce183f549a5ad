"""CPU: the C-ABI library loads, exports every symbol include/uvio_klt.h declares, and refuses to run without a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    inc = os.path.join(ROOT, "include")
    for h in sorted(os.listdir(inc)):
        src = open(os.path.join(inc, h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b(uvio_(?:klt|featdb|tri|desc)_\w+)\s*\(", src))
    return sorted(names)


def test_library_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(built_lib)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/uvio_klt.h but not exported"
    lib.uvio_klt_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.uvio_klt_version()


def test_every_entry_point_cites_the_reference():
    src = open(os.path.join(ROOT, "include", "uvio_klt.h")).read()
    for ref in ("TrackKLT.cpp:34-94", "TrackBase.h:134-146", "TrackBase.cpp:230-248", "FeatureDatabase.cpp:59", "sensor_data.h:55-80",
                "Grider_GRID.h:125", "Grider_GRID.h:163-174", "TrackKLT.cpp:671,858", "TrackKLT.cpp:860-873"):
        assert ref in src, ref


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from uvio_b200 import configs
    from uvio_b200.track import UvioError, make_tracker
    with pytest.raises(UvioError) as e:
        make_tracker(configs.EUROC_MAV)
    assert e.value.code == 3            # UVIO_E_CUDA: no device, no tracker


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "uvio_b200")
    for dirpath, _d, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".h")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in txt and "from oracle" not in txt and "oracle/" not in txt.replace("no oracle/", ""), f
