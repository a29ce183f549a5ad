"""CPU: the UVIO_USE_REAL_OV_CORE branch of the C++ shim compiles (syntax-only) against the reference's own TrackBase.h /
CamBase.h / CamEqui.h / FeatureDatabase.h / sensor_data.h, with stub OpenCV / Eigen / Boost headers.  Skipped where the
reference tree is absent (the GPU box)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/ov_core/src"


@pytest.mark.skipif(not os.path.isdir(REF), reason="the reference tree is not present on this machine")
def test_shim_compiles_against_the_reference_headers():
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-I", os.path.join(ROOT, "tests", "cpp", "stubs"), "-I", REF, "-I", ROOT,
           os.path.join(ROOT, "tests", "cpp", "real_header_check.cpp")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    assert res.returncode == 0, res.stderr


@pytest.mark.skipif(not os.path.isdir(REF), reason="the reference tree is not present on this machine")
def test_an_override_of_a_plain_trackbase_member_would_not_compile(tmp_path):
    """The check has teeth: declaring get_last_ids() `override` (what round 1's shim did) is rejected by the real header."""
    src = tmp_path / "bad.cpp"
    src.write_text('#define UVIO_USE_REAL_OV_CORE\n#include "uvio_b200/cpp/TrackKLTGpu.h"\n'
                   "struct Bad : ov_core::TrackKLTGpu { using TrackKLTGpu::TrackKLTGpu;\n"
                   "  std::unordered_map<size_t, std::vector<size_t>> get_last_ids() override { return {}; } };\nint main() { return 0; }\n")
    res = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "tests", "cpp", "stubs"), "-I", REF, "-I", ROOT,
                          str(src)], capture_output=True, text=True)
    assert res.returncode != 0 and "override" in res.stderr
