"""Build recipe of libStarFlashAttention.so (nvcc, sm_100a only, in-tree output).

    python starflashattention_b200/build.py [--force] [--verbose]

The library is written to ``starflashattention_b200/lib/libStarFlashAttention.so`` so that it travels
with the repository snapshot to the GPU box.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import argparse
import concurrent.futures
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(PKG, "lib", "obj")
LIB = os.path.join(PKG, "lib", "libStarFlashAttention.so")
# Test / tuning build of the same sources with -DSFA_PROBES: extra prefill instantiations selectable through
# SFA_FWD_* environment variables and the event trace.  Never loaded by the product path (SFA_PROBES_LIB=1 selects it).
OBJ_PROBES = os.path.join(PKG, "lib", "obj_probes")
LIB_PROBES = os.path.join(PKG, "lib", "libStarFlashAttention_probes.so")

SOURCES = ["common.cu", "decode_sm100.cu", "rope_cache.cu", "shard_sm100.cu", "fwd_sm100.cu", "host_pipeline.cu", "capi.cu", "decode_static.cu"]
# the probe library also carries the round-1 prefill kernel (one CTA per query block) for A/B runs: SFA_FWD_V1=1
PROBE_SOURCES = SOURCES + ["fwd_sm100_v1.cu"]
HEADERS = [os.path.join(CSRC, "internal.h"), os.path.join(CSRC, "sm100_ptx.cuh"), os.path.join(CSRC, "rope.cuh"), os.path.join(CSRC, "merge.cuh"),
           os.path.join(ROOT, "include", "star_flash_attn.h"), os.path.join(ROOT, "include", "src", "flash_attn.h"),
           os.path.join(ROOT, "include", "src", "params.h"), os.path.join(ROOT, "include", "src", "traits.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; the library cannot be built")
    return exe


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _compile(src: str, verbose: bool, probes: bool = False) -> str:
    obj = os.path.join(OBJ_PROBES if probes else OBJ, src.replace(".cu", ".o"))
    cmd = [nvcc(), *NVCC_FLAGS, *(["-DSFA_PROBES"] if probes else []), "-c", os.path.join(CSRC, src), "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        sys.stderr.write(r.stderr)
    return obj


def build(force: bool = False, verbose: bool = False, probes: bool = False) -> str:
    obj_dir, lib = (OBJ_PROBES, LIB_PROBES) if probes else (OBJ, LIB)
    os.makedirs(obj_dir, exist_ok=True)
    sources = PROBE_SOURCES if probes else SOURCES
    todo = [s for s in sources
            if force or _stale(os.path.join(obj_dir, s.replace(".cu", ".o")), [os.path.join(CSRC, s), *HEADERS])]
    if todo:
        with concurrent.futures.ThreadPoolExecutor(max_workers=len(todo)) as ex:
            list(ex.map(lambda s: _compile(s, verbose, probes), todo))
    objs = [os.path.join(obj_dir, s.replace(".cu", ".o")) for s in sources]
    if force or todo or _stale(lib, objs):
        cmd = [nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib, *objs]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return lib


def torch_ext_path() -> str:
    import sysconfig
    return os.path.join(ROOT, "star_flash_attn" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_torch_ext(force: bool = False) -> str:
    """The reference-facing `star_flash_attn` PyTorch extension (csrc/flash_api.cpp, host-only C++) linked against
    libStarFlashAttention.so.  Written next to the package at the repository root so `import star_flash_attn`
    works from there, as it does after the reference's `setup.py install` (reference setup.py:9-30)."""
    import sysconfig
    from torch.utils import cpp_extension as ce
    import torch
    lib = build()
    out = torch_ext_path()
    src = os.path.join(CSRC, "flash_api.cpp")
    if not (force or _stale(out, [src, lib, os.path.join(ROOT, "include", "star_flash_attn.h")])):
        return out
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    inc = [f"-I{d}" for d in ce.include_paths(device_type="cuda")] + [f"-I{sysconfig.get_paths()['include']}",
                                                                  f"-I{os.path.join(ROOT, 'include')}"]
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=star_flash_attn",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch.compiled_with_cxx11_abi())}",
           *inc, src, "-o", out, f"-L{tlib}", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
           "-ltorch_python", f"-L{os.path.dirname(lib)}", "-lStarFlashAttention",
           f"-Wl,-rpath,{tlib}", "-Wl,-rpath,$ORIGIN/starflashattention_b200/lib"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"star_flash_attn extension build failed:\n{r.stdout}\n{r.stderr}")
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--torch-ext", action="store_true", help="also build the star_flash_attn PyTorch extension")
    ap.add_argument("--probes", action="store_true", help="also build the -DSFA_PROBES tuning library")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
    if a.probes:
        print(build(a.force, a.verbose, probes=True))
    if a.torch_ext:
        print(build_torch_ext(a.force))
