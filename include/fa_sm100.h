/*
 * fa_sm100.h — C ABI of the B200 (sm_100a) FlashAttention-2 forward operator.
 *
 * This is the drop-in boundary for the reference's hot path.  It replaces, one for one,
 *
 *   torch::Tensor flash_attention_fwd(Tensor& q, Tensor& k, Tensor& v, float softmax_scale, bool causal)
 *       -- reference csrc/flash_attention_api.cpp:14-135 (pybind at :137-141), and below it
 *   struct FlashAttentionParams            -- reference csrc/flash_attention.h:5-37
 *   run_flash_attention<T,kHeaddim,Causal> -- reference csrc/flash_attention.h:40-41,
 *                                             csrc/flash_attention_impl.cu:30-49
 *   type/bool/headdim dispatchers          -- reference csrc/kernel_dispatcher.h:20-52
 *
 * with plain pointers and sizes: no torch types, no C++ in the signatures.  The caller owns
 * q, k, v and o (o must be allocated by the caller with the shape of q; the reference uses
 * empty_like(q), so o normally has q's strides — any strides with unit last-dim stride are
 * accepted).  All calls are asynchronous on `stream`, never synchronise the device, never
 * allocate, and never call exit().  Errors come back as a non-zero status plus a thread-local
 * message from fa_sm100_last_error().
 *
 * Semantics (identical to the reference):
 *   layout  q [B, Hq, Sq, D], k/v [B, Hkv, Sk, D], o [B, Hq, Sq, D]; strides in ELEMENTS,
 *           last-dim stride must be 1 (csrc/flash_attention_api.cpp:44-49)
 *   dtype   fp16 or bf16, fp32 accumulation, P rounded to the input dtype before P·V
 *   GQA     q-head h reads kv-head h / (Hq/Hkv) (csrc/flash_attention_template.cuh:157-160)
 *   causal  bottom-right aligned: column n is visible to row m iff n <= m + (Sk - Sq)
 *           (csrc/mask.cuh:37-39)
 *   scale   `softmax_scale` is the natural-base scale (e.g. D^-0.5); log2(e) is folded inside
 *           (the reference folds it at csrc/flash_attention_api.cpp:87)
 *   D       multiple of 8, <= 128 (csrc/flash_attention_api.cpp:52-55)
 *   Sq == 1 the reference packs the q heads of a group and drops `causal`
 *           (csrc/flash_attention_api.cpp:72-83); with bottom-right alignment a single query row
 *           sees every key, so the result is the same and no special case is needed here.
 * Difference, documented: a query row with no visible key (only possible when causal and
 * Sq > Sk) yields zeros here; the reference yields zeros or an unnormalised mean of V depending
 * on tile skipping.
 */
#ifndef FA_SM100_H_
#define FA_SM100_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum fa_sm100_dtype { FA_SM100_FP16 = 0, FA_SM100_BF16 = 1 };

enum fa_sm100_status {
    FA_SM100_OK = 0,
    FA_SM100_ERR_INVALID_ARGUMENT = 1, /* shape / stride / dtype / alignment check failed   */
    FA_SM100_ERR_UNSUPPORTED_DEVICE = 2, /* not an sm_100 GPU                                */
    FA_SM100_ERR_CUDA = 3,             /* CUDA runtime / driver error (message has detail) */
    FA_SM100_ERR_KERNEL_ABORT = 4      /* in-kernel watchdog fired (fa_sm100_check_device) */
};

/* Flags for fa_sm100_fwd_params.flags */
#define FA_SM100_FLAG_NO_PACK_GQA 1u   /* do not pack the q heads of a kv group into one M tile */
#define FA_SM100_FLAG_FORCE_PACK_GQA 2u /* pack whenever Hq/Hkv > 1 (default: heuristic)        */
#define FA_SM100_FLAG_KERNEL_V1 4u      /* debug: first-generation (non-persistent) kernel       */
#define FA_SM100_FLAG_NO_DECODE 8u      /* operator-level routing hint: keep seqlen_q == 1 on
                                           fa_sm100_fwd instead of fa_sm100_fwd_decode            */
/* Tuning bits (honoured by FA_SM100_TUNE builds, ignored otherwise):
 *   bits 8-11  1 + number of exp2 evaluations per 4 element pairs moved to the FMA pipe (0 = default)
 *   bits 12-13 P hand-off split: 0 = default, 1 = 64 keys first, 2 = 96, 3 = no split
 *   bits 14-15 softmax threads per query row: 0 = default, 1 or 2                                 */
#define FA_SM100_FLAG_TUNE_EMU(n) ((uint32_t)((n) + 1) << 8)
#define FA_SM100_FLAG_TUNE_SPLIT(sel) ((uint32_t)(sel) << 12)
#define FA_SM100_FLAG_TUNE_HALVES(n) ((uint32_t)(n) << 14)

typedef struct fa_sm100_fwd_params {
    const void* q;      /* device pointers, 16-byte aligned */
    const void* k;
    const void* v;
    void* o;
    int64_t batch;
    int64_t heads_q;
    int64_t heads_kv;
    int64_t seqlen_q;
    int64_t seqlen_kv;
    int64_t headdim;
    /* element strides: [0] batch, [1] head, [2] sequence; last-dim stride is 1 */
    int64_t q_strides[3];
    int64_t k_strides[3];
    int64_t v_strides[3];
    int64_t o_strides[3];
    float softmax_scale; /* natural scale; the kernel works in the exp2 domain internally */
    int32_t causal;      /* 0 / 1 */
    int32_t dtype;       /* enum fa_sm100_dtype */
    int32_t device;      /* CUDA device ordinal owning the pointers */
    uint32_t flags;      /* FA_SM100_FLAG_* */
} fa_sm100_fwd_params;

/* Launch the forward kernel on `stream` (a cudaStream_t passed as void*; NULL = default stream).
 * Returns fa_sm100_status. */
int fa_sm100_fwd(const fa_sm100_fwd_params* params, void* stream);

/* ---- Decode (seqlen_q == 1) with split-KV --------------------------------------------------------
 * Replaces the reference's Sq == 1 branch, which packs the q heads of a kv group into one M tile
 * and runs the prefill kernel with one CTA per (batch, kv-head)
 * (csrc/flash_attention_api.cpp:72-83, 128-133; "Flash Decoding" is a TODO at README.md:20).
 * Same params struct and semantics as fa_sm100_fwd (causal is irrelevant: with bottom-right
 * alignment the single query row sees every key); requires seqlen_q == 1, Hq/Hkv <= 16 and
 * softmax_scale > 0 — anything else is FA_SM100_ERR_INVALID_ARGUMENT and belongs to fa_sm100_fwd.
 * The K/V sequence of each (batch, kv-head) unit is cut into `nsplit` chunks, one CTA each; the
 * last CTA of a unit merges the partial (max, sum, O) triples, so it is still ONE launch.
 *
 * fa_sm100_decode_plan reports the split the library would use and the workspace it needs
 * (0 bytes when nsplit == 1).  The caller owns the workspace: 256-byte aligned device memory whose
 * first ceil(4·B·Hkv / 256)·256 bytes are ZERO before the first launch that uses it; every launch
 * leaves them zero again, so one buffer serves all stream-ordered launches. */
int fa_sm100_decode_plan(const fa_sm100_fwd_params* params, int32_t* nsplit, int32_t* chunk_keys,
                         int64_t* workspace_bytes);
int fa_sm100_fwd_decode(const fa_sm100_fwd_params* params, void* workspace, int64_t workspace_bytes,
                        void* stream);

/* ---- Variable-length (packed) batches ---------------------------------------------------------------
 * SURVEY §8(f) rank 4; the reference lists "varlen" as a TODO (README.md:18) and has no code for it,
 * so the semantics follow its fixed-length path sequence by sequence: sequence i owns the token rows
 * cu_seqlens_q[i] .. cu_seqlens_q[i+1] of q / o ([total_q, Hq, D]) and cu_seqlens_kv[i] ..
 * cu_seqlens_kv[i+1] of k / v ([total_kv, Hkv, D]) and is computed exactly like one batch entry of
 * fa_sm100_fwd (same GQA rule, same bottom-right causal alignment per sequence).
 * cu_seqlens_* are HOST arrays of num_seqs + 1 non-decreasing int32 offsets starting at 0 (the
 * caller already has them on the host to size its buffers; no device read, no sync).  First
 * implementation: one launch of the forward kernel per non-empty sequence on `stream`, each with its
 * own tensor maps clipped to the sequence (so no row of a neighbouring sequence is ever read or
 * written).  Sequences with no queries are skipped; a sequence with queries and no keys writes
 * zeros.  Token strides must keep every sequence base 16-byte aligned (token stride % 8 == 0). */
typedef struct fa_sm100_varlen_params {
    const void* q; /* [total_q,  Hq,  D] */
    const void* k; /* [total_kv, Hkv, D] */
    const void* v;
    void* o;       /* [total_q,  Hq,  D] */
    const int32_t* cu_seqlens_q;  /* host, num_seqs + 1 entries */
    const int32_t* cu_seqlens_kv; /* host, num_seqs + 1 entries */
    int64_t num_seqs;
    int64_t heads_q;
    int64_t heads_kv;
    int64_t headdim;
    /* element strides: [0] token, [1] head; last-dim stride is 1 */
    int64_t q_strides[2];
    int64_t k_strides[2];
    int64_t v_strides[2];
    int64_t o_strides[2];
    float softmax_scale;
    int32_t causal;
    int32_t dtype;
    int32_t device;
    uint32_t flags;
} fa_sm100_varlen_params;

int fa_sm100_fwd_varlen(const fa_sm100_varlen_params* params, void* stream);

/* ---- Rotary position embedding on q and k, one pass ----------------------------------------------
 * Replaces apply_rotary_pos_emb / rotate_half of the reference's model path
 * (models/rope_attn_fwd.py:8-12, 14-38, call at :87-88):
 *     out = x * cos + rotate_half(x) * sin,   rotate_half(x) = cat(-x[D/2:], x[:D/2])
 * for q [B, Hq, S, D] and k [B, Hkv, S, D] (any strides with unit last-dim stride — the
 * [B, S, H, D]-memory views of the projections run as they are), cos / sin [B or 1, S, D] of the
 * same dtype.  Rounding follows PyTorch's 16-bit evaluation exactly (each multiply and the add
 * rounded to the dtype), so the result is bit-equal to the reference expression.  q_out / k_out
 * may alias q / k (in place).  D must be a multiple of 16. */
typedef struct fa_sm100_rope_params {
    const void* q;
    const void* k;
    void* q_out;
    void* k_out;
    const void* cos;
    const void* sin;
    int64_t batch;
    int64_t heads_q;
    int64_t heads_kv;
    int64_t seqlen;
    int64_t headdim;
    /* element strides: [0] batch, [1] head, [2] sequence; last-dim stride is 1 */
    int64_t q_strides[3];
    int64_t k_strides[3];
    int64_t q_out_strides[3];
    int64_t k_out_strides[3];
    int64_t cos_sin_strides[2]; /* [0] batch (ignored when batch == 1; 0 = broadcast), [1] sequence */
    int32_t dtype;              /* enum fa_sm100_dtype */
    int32_t device;
} fa_sm100_rope_params;

int fa_sm100_rope(const fa_sm100_rope_params* params, void* stream);

/* Thread-local, NUL-terminated description of the last non-zero status on this thread. */
const char* fa_sm100_last_error(void);

/* Synchronises `device` and reports whether any kernel launched by this library since the last
 * call hit its in-kernel watchdog (test / debug aid; not used on the hot path). */
int fa_sm100_check_device(int device);

/* Library / kernel introspection used by tests and bench: number of kernel launches issued by
 * this library in this process, and a static description of the build. */
uint64_t fa_sm100_launch_count(void);
const char* fa_sm100_build_info(void);

/* Describe the variant fa_sm100_fwd would launch for `params` (no launch): fills grid size,
 * dynamic shared memory bytes, pack factor (1 = unpacked) and q rows per head per tile. */
int fa_sm100_plan(const fa_sm100_fwd_params* params, int32_t* grid, int32_t* smem_bytes,
                  int32_t* pack_g, int32_t* rows_per_head);

/* UMMA/TMA primitive self-test (tests/test_primitives.py): one CTA computes S = Q·Kᵀ
 * (128 x 128, fp32) and O = round(S * p_scale)·V (128 x D, fp32) with the same descriptors the
 * forward kernel uses.  q, k, v: [128, D] row-major device buffers of `dtype`;
 * s_out: [128,128] fp32; o_out: [128, D] fp32.  v_lbo / v_sbo override the V descriptor's
 * byte offsets when non-zero. */
int fa_sm100_selftest_umma(const void* q, const void* k, const void* v, float* s_out,
                           float* o_out, int32_t headdim, int32_t dtype, float p_scale,
                           uint32_t v_lbo, uint32_t v_sbo, int32_t device, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FA_SM100_H_ */
