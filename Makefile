# Builds build/libStarFlashAttention.so (sm_100a only) and the C++ example, the layout the reference's
# Makefile produces (reference Makefile:1-36; retargeted, NVTX3 header-only, no cuBLAS).
PY ?= python
NVCC ?= nvcc
ARCH = -gencode arch=compute_100a,code=sm_100a
PKG_LIB = starflashattention_b200/lib/libStarFlashAttention.so

all: lib example

lib:
	$(PY) starflashattention_b200/build.py
	@mkdir -p build
	@cp -f $(PKG_LIB) build/libStarFlashAttention.so

example: lib
	$(NVCC) -o build/testFlashDecoder examples/cpp/testFlashDecoder.cc -O2 -std=c++17 $(ARCH) -lineinfo \
	  -I include -L build -lStarFlashAttention -Xlinker -rpath -Xlinker '$$ORIGIN'

# C++ callers of the MANGLED entry points (run by tests/test_capi_gpu.py on the GPU box)
cpptest: lib
	$(NVCC) -o build/test_mangled_entry tests/cpp/test_mangled_entry.cc -O2 -std=c++17 $(ARCH) -lineinfo \
	  -I include -L build -lStarFlashAttention -Xlinker -rpath -Xlinker '$$ORIGIN'

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf build starflashattention_b200/lib

.PHONY: all lib example cpptest oracle clean
