import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `pytest -m gpu`)")


def _have_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built_lib():
    from uvio_b200 import build
    return build.build()


# ---- host simulation (test scaffolding): the engine + ABI over CPU stand-ins of the launch wrappers ----------------------------
HERE = os.path.dirname(os.path.abspath(__file__))
SIM = os.path.join(HERE, "hostsim")
SIM_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-Wno-unknown-pragmas"]


def SIM_SOURCES():
    """engine + C ABIs over the CPU stand-ins of every launch wrapper (tests/hostsim)"""
    csrc = os.path.join(ROOT, "uvio_b200", "csrc")
    return ([os.path.join(csrc, f) for f in ("engine.cpp", "engine_upload.cpp", "engine_host.cpp", "engine_device.cpp", "engine_stages.cpp", "abi.cpp", "featdb.cpp", "desc_engine.cpp")] +
            [os.path.join(SIM, f) for f in ("launch_sim.cpp", "featdb_sim.cpp", "desc_sim.cpp")])


def sim_build(out, srcs, extra=()):
    import subprocess
    if os.environ.get("UVIO_SIM_SANITIZE"):      # tools/asan_sim.sh: the same bodies under AddressSanitizer + UBSan
        out = out.replace(os.sep + "_sim" + os.sep, os.sep + "_sim_asan" + os.sep)
        extra = list(extra) + ["-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-g", "-O1", "-fno-sanitize-recover=undefined"]
    os.makedirs(os.path.dirname(out), exist_ok=True)
    deps = srcs + [os.path.join(ROOT, "uvio_b200", "csrc", f) for f in os.listdir(os.path.join(ROOT, "uvio_b200", "csrc"))]
    deps += [os.path.join(SIM, f) for f in os.listdir(SIM) if f.endswith((".cpp", ".h"))]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        subprocess.check_call(["g++"] + SIM_FLAGS + list(extra) + ["-o", out] + srcs + ["-lpthread"])
    return out


@pytest.fixture()
def sim_tracker(monkeypatch):
    """make(cfg, **kw) -> tracker running engine.cpp + abi.cpp over tests/hostsim/launch_sim.cpp (no GPU, never the product)."""
    from uvio_b200 import track
    csrc = os.path.join(ROOT, "uvio_b200", "csrc")
    lib = sim_build(os.path.join(SIM, "_sim", "libuvio_klt_sim.so"),
                    SIM_SOURCES(), ["-I", SIM])
    monkeypatch.setattr(track, "LIB_PATH", lib)
    monkeypatch.setattr(track, "_lib", None)
    yield lambda cfg, **kw: track.make_tracker(cfg, **kw)
    monkeypatch.setattr(track, "_lib", None)
