// Compile-time shim kept so that `#include <src/traits.h>` and Traits<half,128,128> still name
// something (reference: src/traits.h:7-26, a shared-memory size calculator).  Tile shapes and
// shared-memory budgets of the sm_100a kernels are chosen at run time by the host launcher
// (launch_fwd in starflashattention_b200/csrc/fwd_sm100.cu, launch_decode in csrc/decode_sm100.cu); nothing in the
// library reads this struct.
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>

template <typename elem_type, int head_dim, int n_elem_per_blockN>
struct Traits
{
    static constexpr int warp_size = 32;
    static constexpr int n_elem_per_vec = 16 / sizeof(elem_type);
    static constexpr int n_elem_per_thread = head_dim / warp_size;
    static constexpr int kHeadDim = head_dim;
    static constexpr int kBlockN = n_elem_per_blockN;
    // The decode kernel streams K/V through registers; it uses no dynamic shared memory.
    static constexpr int smemSize = 0;
};
