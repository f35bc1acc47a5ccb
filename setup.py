"""Builds the reference-facing PyTorch extension module `star_flash_attn` (pybind11 over the C ABI) and ships the
package `starflashattention_b200`: the working form of the reference's setup.py (reference setup.py:1-30, which
forgets `import os` (:5) and links a prebuilt build/libStarFlashAttention.so by absolute path).

    python setup.py build_ext --inplace     # star_flash_attn.*.so next to this file
    pip install --no-build-isolation .      # package + extension module

The CUDA library itself (sm_100a only) is compiled first by starflashattention_b200/build.py (nvcc, in-tree);
the extension is host-only C++ (csrc/flash_api.cpp), so a plain CppExtension linked against that library suffices.
"""
import importlib.util
import os

from setuptools import setup
from torch.utils import cpp_extension

this_dir = os.path.dirname(os.path.abspath(__file__))

# build.py is loaded by path: importing the package would dlopen a library that may not exist yet
_spec = importlib.util.spec_from_file_location("_sfa_build", os.path.join(this_dir, "starflashattention_b200", "build.py"))
_build = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_build)
lib_path = _build.build()
lib_dir = os.path.dirname(lib_path)

setup(
    name="star_flash_attn",
    version="0.2.0",
    description="B200-native attention forward / split-KV decode behind StarFlashAttention's extension interface",
    packages=["starflashattention_b200"],
    package_data={"starflashattention_b200": ["lib/libStarFlashAttention.so"]},
    ext_modules=[
        cpp_extension.CppExtension(
            name="star_flash_attn",
            sources=[os.path.join("starflashattention_b200", "csrc", "flash_api.cpp")],
            include_dirs=[os.path.join(this_dir, "include"), *cpp_extension.include_paths(device_type="cuda")],
            library_dirs=[lib_dir, *cpp_extension.library_paths(device_type="cuda")],
            libraries=["StarFlashAttention", "c10_cuda", "torch_cuda"],
            # found next to the module when built in place, and inside the installed package
            runtime_library_dirs=[lib_dir, "$ORIGIN/starflashattention_b200/lib"],
            extra_compile_args=["-O2", "-std=c++17"],
        )
    ],
    cmdclass={"build_ext": cpp_extension.BuildExtension},
)
