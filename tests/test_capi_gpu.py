"""GPU: the C++-MANGLED entry points, called from C++ binaries as a C++ user of the reference calls them
(/root/reference examples/cpp/testFlashDecoder.cc:97-103 -> run_flash_decoder<half>; plus the typed additions
run_flash_fwd<T> / run_flash_decode<T>).  tests/cpp/test_mangled_entry.cc compares each with the C-ABI entry point
bit for bit (the C ABI itself is what the other GPU tests check against the oracle) and runs the reference
example's all-ones fixture; examples/cpp/testFlashDecoder.cc is the reference example's six shapes."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _binary(name, target):
    path = os.path.join(ROOT, "build", name)
    lib = os.path.join(ROOT, "build", "libStarFlashAttention.so")
    pkg_lib = os.path.join(ROOT, "starflashattention_b200", "lib", "libStarFlashAttention.so")
    stale = not os.path.exists(path) or not os.path.exists(lib) or os.path.getmtime(lib) < os.path.getmtime(pkg_lib)
    if stale:
        subprocess.run(["make", "-C", ROOT, "-s", target], check=True)
    return path


def test_mangled_entry_points_match_c_abi(built_library):
    exe = _binary("test_mangled_entry", "cpptest")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all mangled entry points match the C ABI" in r.stdout
    for what in ("run_flash_decoder<half> O", "run_flash_fwd<half> O", "run_flash_fwd<bf16> O",
                 "run_flash_decode<half> O", "run_flash_decode<bf16> O", "run_flash_decoder<half> K cache after the append"):
        assert f"[equal] {what}" in r.stdout, r.stdout


def test_reference_cpp_example_all_ones(built_library):
    """The reference example's own fixture (six shapes, V == 1 => O == 1) through run_flash_decoder<half>."""
    exe = _binary("testFlashDecoder", "example")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count(" ok") == 6 and "MISMATCH" not in r.stdout, r.stdout
