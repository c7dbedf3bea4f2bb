// TEST INFRASTRUCTURE ONLY — never imported by the product path.
//
// C-ABI launcher around the *unmodified* reference kernel translation unit
// (/root/reference/csrc/flash_attention_impl.cu, compiled where it lies by
// oracle/build_ref.sh into oracle/_ref/libref_fa.so).  It fills the reference's
// own FlashAttentionParams (csrc/flash_attention.h:5-37) the way the reference
// host function does (csrc/flash_attention_api.cpp:87-113: log2(e) folded into
// the scale, element strides) and calls run_flash_attention<T,128,IsCausal>
// (csrc/flash_attention_impl.cu:30-49).  The torch-dependent validation and the
// Sq==1 packing reshape of flash_attention_api.cpp:72-83 are done by the Python
// wrapper oracle/ref_kernel.py, because this shim must not link torch.
#include <cuda_runtime.h>
#include <cstdint>
#include <cmath>
#include <cutlass/numeric_types.h>
#include "flash_attention.h"

extern "C" int ref_fa_fwd(void* q, void* k, void* v, void* o,
                          int64_t batch, int64_t heads_q, int64_t heads_kv,
                          int64_t seqlen_q, int64_t seqlen_kv, int64_t headdim,
                          int64_t head_q_per_group,
                          const int64_t* q_strides,  // batch, head, seq (elements)
                          const int64_t* k_strides, const int64_t* v_strides,
                          const int64_t* o_strides,
                          float softmax_scale, int causal, int is_bf16, void* stream) {
    flash_attention::FlashAttentionParams p = {
        q, k, v, o,
        batch, heads_q, heads_kv, seqlen_q, seqlen_kv, headdim, head_q_per_group,
        q_strides[0], k_strides[0], v_strides[0], o_strides[0],
        q_strides[1], k_strides[1], v_strides[1], o_strides[1],
        q_strides[2], k_strides[2], v_strides[2], o_strides[2],
        softmax_scale * float(M_LOG2E)};
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (headdim > 128) return 1;
    if (is_bf16) {
        if (causal) flash_attention::run_flash_attention<cutlass::bfloat16_t, 128, true>(p, s);
        else        flash_attention::run_flash_attention<cutlass::bfloat16_t, 128, false>(p, s);
    } else {
        if (causal) flash_attention::run_flash_attention<cutlass::half_t, 128, true>(p, s);
        else        flash_attention::run_flash_attention<cutlass::half_t, 128, false>(p, s);
    }
    return int(cudaGetLastError());
}
