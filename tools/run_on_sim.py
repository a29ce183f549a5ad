"""Development aid: runs GPU test functions against the host-simulation build of the library (tests/hostsim), e.g.
    python tools/run_on_sim.py tests/test_gpu_edge.py test_few_features_match_oracle 12
so host logic and kernel bodies can be debugged without a GPU.  Never used by the product, the bench or the test suite."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import conftest  # noqa: E402
from uvio_b200 import track  # noqa: E402

csrc = os.path.join(ROOT, "uvio_b200", "csrc")
track.LIB_PATH = conftest.sim_build(os.path.join(conftest.SIM, "_sim", "libuvio_klt_sim.so"),
                                    conftest.SIM_SOURCES(), ["-I", conftest.SIM])
mod = importlib.import_module(os.path.basename(sys.argv[1])[:-3])
fn = getattr(mod, sys.argv[2])
args = [int(a) if a.lstrip("-").isdigit() else a for a in sys.argv[3:]]
print(fn(None, *args) if "built_lib" in fn.__code__.co_varnames[:1] else fn(*args))
