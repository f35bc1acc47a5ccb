"""CPU: the C-ABI library loads, exports every symbol include/star_flash_attn.h declares, keeps the
reference's C++-mangled drop-ins, validates arguments, and refuses to compute without a B200."""
import ctypes
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    text = open(os.path.join(ROOT, "include", "star_flash_attn.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sfa_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_are_exported(built_library):
    from starflashattention_b200 import _capi
    declared = _declared_functions()
    assert len(declared) >= 12
    assert set(declared) == set(_capi.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(_capi.lib, name), name


def test_reference_mangled_symbols_present(built_library):
    from starflashattention_b200 import _capi
    out = subprocess.run(["nm", "-D", "--defined-only", built_library], capture_output=True, text=True).stdout
    for sym in _capi.MANGLED_DROP_INS:
        assert sym in out, sym
    # typed additions
    assert "run_flash_fwd" in out and "run_flash_decode" in out


def test_abi_version_and_struct_sizes(built_library):
    from starflashattention_b200 import _capi
    assert _capi.lib.sfa_abi_version() == 2
    # offsets implied by include/star_flash_attn.h (LP64): 5 pointers, 6 int32, 12 int64, float + 2 int32 (+pad)
    assert ctypes.sizeof(_capi.FwdArgs) == 5 * 8 + 6 * 4 + 12 * 8 + 4 + 2 * 4 + 4
    assert ctypes.sizeof(_capi.DecodeArgs) == 6 * 8 + 5 * 4 + 4 + 10 * 8 + 3 * 4 + 4 + 8 + 8
    assert ctypes.sizeof(_capi.DecodePagedArgs) == ctypes.sizeof(_capi.DecodeArgs) + 8 + 8 + 4 + 4
    assert ctypes.sizeof(_capi.DecoderStepArgs) == 8 * 8 + 4 * 4 + 4 + 6 * 4 + 4 + 8 + 8     # + workspace, workspace_bytes
    assert ctypes.sizeof(_capi.FwdHostArgs) == ctypes.sizeof(_capi.FwdArgs) + 8
    assert ctypes.sizeof(_capi.DecodeHostArgs) == ctypes.sizeof(_capi.DecodeArgs) + 8


def test_argument_validation_without_gpu(built_library):
    from starflashattention_b200 import _capi
    lib = _capi.lib
    assert lib.sfa_fwd(None, None) == _capi.SFA_ERR_INVALID
    assert b"null" in lib.sfa_last_error()
    a = _capi.FwdArgs()
    assert lib.sfa_fwd(a, None) == _capi.SFA_ERR_INVALID          # null tensors
    buf = (ctypes.c_char * 4096)()
    p = (ctypes.addressof(buf) + 15) // 16 * 16
    a.q = a.k = a.v = a.o = p
    a.batch, a.seqlen_q, a.seqlen_k, a.heads_q, a.heads_kv, a.head_dim = 1, 4, 4, 3, 2, 128
    assert lib.sfa_fwd(a, None) == _capi.SFA_ERR_INVALID          # heads_q % heads_kv
    a.heads_q = 2
    a.head_dim = 96
    assert lib.sfa_fwd(a, None) == _capi.SFA_ERR_UNSUPPORTED
    a.head_dim = 128
    a.q_row_stride = 100                                          # not a multiple of 8
    assert lib.sfa_fwd(a, None) == _capi.SFA_ERR_INVALID
    d = _capi.DecodeArgs()
    d.dtype = 7
    assert lib.sfa_decode(d, None) == _capi.SFA_ERR_INVALID
    assert lib.sfa_decode_workspace_bytes(8, 32, 128, 1) == 0
    pg = _capi.DecodePagedArgs()
    assert lib.sfa_decode_paged(None, None) == _capi.SFA_ERR_INVALID
    assert lib.sfa_decode_paged(pg, None) == _capi.SFA_ERR_INVALID       # null block table
    pg.block_table, pg.num_pages, pg.block_table_stride = p, 4, 1
    for bad in (0, 8, 24, 48):
        pg.page_size = bad
        assert lib.sfa_decode_paged(pg, None) == _capi.SFA_ERR_INVALID and b"page_size" in lib.sfa_last_error()
    pg.page_size, pg.base.seqlen_k = 16, 64                              # 4 blocks needed, rows hold 1
    assert lib.sfa_decode_paged(pg, None) == _capi.SFA_ERR_INVALID and b"block table" in lib.sfa_last_error()
    assert lib.sfa_decode_workspace_bytes(8, 32, 128, 16) >= 8 * 32 * 16 * (128 * 4 + 8)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(built_library):
    from starflashattention_b200 import _capi, interface
    lib = _capi.lib
    buf = (ctypes.c_char * 4096)()
    p = (ctypes.addressof(buf) + 15) // 16 * 16
    assert lib.sfa_rotary_table(p, p, 4, 8, _capi.SFA_FP16, None) == _capi.SFA_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.sfa_last_error()
    q = torch.zeros(1, 4, 1, 64, dtype=torch.float16)
    with pytest.raises(ValueError, match="no CPU fallback"):
        interface.fwd(q, q, q)


def test_product_code_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "starflashattention_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("the oracle", ""), os.path.join(dirpath, f)


def test_torch_extension_keeps_the_reference_signature(torch_ext):
    """`star_flash_attn.mha_fwd_cuda` takes the reference's 16 keyword arguments (src/flash_api.cpp:74-79)."""
    doc = torch_ext.mha_fwd_cuda.__doc__
    names = ["qkv", "q_bias", "k_bias", "v_bias", "k_cache_table", "v_cache_table", "seq_len", "o", "batch_size",
             "memory_max_len", "num_heads", "head_dim", "rotary_embedding_dim", "max_input_length", "num_layer", "idx_layer"]
    pos = [doc.index(n + ":") for n in names]
    assert pos == sorted(pos)
    assert hasattr(torch_ext, "fwd") and hasattr(torch_ext, "decode")
    t = torch.zeros(1, 3, 1, 128, dtype=torch.float16)
    with pytest.raises(RuntimeError, match="CUDA"):
        torch_ext.mha_fwd_cuda(t, t, t, t, t, t, torch.zeros(1, dtype=torch.int32), t, 1, 8, 1, 128, 128, 8, 1, 0)


def test_setup_py_build_ext(tmp_path):
    """`python setup.py build_ext` (the fixed form of the reference's setup.py:1-30) compiles csrc/flash_api.cpp and
    links it against libStarFlashAttention.so; the module must import and expose mha_fwd_cuda."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "setup.py", "-q", "build_ext", "--build-lib", str(tmp_path), "--build-temp",
                        str(tmp_path / "tmp")], cwd=root, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    built = [f for f in os.listdir(tmp_path) if f.startswith("star_flash_attn") and f.endswith(".so")]
    assert built, os.listdir(tmp_path)
    code = ("import sys, torch; sys.path.insert(0, %r); import star_flash_attn as m; "
            "assert hasattr(m, 'mha_fwd_cuda'); print('ok', m.__file__)" % str(tmp_path))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=str(tmp_path))
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr
