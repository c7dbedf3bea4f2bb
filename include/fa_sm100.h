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
#define FA_SM100_FLAG_NO_DECODE 8u      /* operator-level routing hint: keep seqlen_q == 1 on
                                           fa_sm100_fwd instead of fa_sm100_fwd_decode            */
#define FA_SM100_FLAG_PAGED_ONE_WAVE 16u /* fa_sm100_fwd_decode_paged: the caller knows that the
                                           sequence lengths are close to max_seqlen_k; split every
                                           sequence alike instead of sharing the flattened key space */
/* Tuning bits (honoured by builds with -DFA_SM100_TUNE, ignored by the product build):
 *   bits 8-11  1 + number of exp2 evaluations per 8 element pairs moved to the FMA pipe (0 = default)
 *   bits 12-13 speculative stale-max softmax: 0 = default, 1 = off, 2 = on
 *   bits 14-15 tcgen05.fence::after_thread_sync in front of P·V: 0 = default, 1 = off, 2 = on     */
#define FA_SM100_FLAG_TUNE_EMU(n) ((uint32_t)((n) + 1) << 8)
#define FA_SM100_FLAG_TUNE_SPEC(sel) ((uint32_t)(sel) << 12)
#define FA_SM100_FLAG_TUNE_FENCE(sel) ((uint32_t)(sel) << 14)

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
 * (0 bytes when nsplit == 1).  The caller owns the workspace: 256-byte aligned device memory that
 * holds the partial results of the launch only — no state survives a launch and nothing has to be
 * zeroed, so one buffer (of the largest size any call needs) serves all stream-ordered launches
 * of any shape.  The per-unit arrival counters live in a library-owned ring.  (With 2 .. 8 splits
 * the CTAs of a unit run as one thread-block cluster and merge through distributed shared memory;
 * the workspace of the plan is then not touched, but the contract stays: pass what the plan asks.) */
int fa_sm100_decode_plan(const fa_sm100_fwd_params* params, int32_t* nsplit, int32_t* chunk_keys,
                         int64_t* workspace_bytes);
int fa_sm100_fwd_decode(const fa_sm100_fwd_params* params, void* workspace, int64_t workspace_bytes,
                        void* stream);

/* ---- Decode (seqlen_q == 1) that allocates nothing -----------------------------------------------
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that
 * forbid it).  Same params, validation, split plan, kernel and results (bit-equal) as
 * fa_sm100_fwd_decode, but the split partials and the per-unit arrival counters live in a
 * file-scope device array of FA_SM100_STATIC_SCRATCH_BYTES inside the library instead of a caller
 * workspace and the library's allocated counter ring.  Neither the call nor its once-per-device
 * set-up (cudaGetSymbolAddress with that device current) calls cudaMalloc, cudaFree or any
 * allocating routine; every GPU that loads the library holds the array until the process exits.
 * A split that needs more than the array holds (fa_sm100_decode_plan's workspace_bytes >
 * FA_SM100_STATIC_SCRATCH_BYTES - 2048) is FA_SM100_ERR_INVALID_ARGUMENT with the sizes in
 * fa_sm100_last_error — with the one-wave plan on a 148-SM B200 the largest request is 2.46 MB,
 * so the limit only binds on parts with more than 251 SMs.  All calls on a device share the
 * array, which keeps the last call's contents: the counters are cleared in stream order before
 * every launch, the partials are written before they are read, so calls ordered on one stream
 * (or by events) are safe — calls that OVERLAP on the GPU (different streams, no ordering) race
 * on it and must use fa_sm100_fwd_decode with separate workspaces.  This entry point does not
 * consult or set the device's sticky watchdog status (the decode kernel has no in-kernel waits). */
#define FA_SM100_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int fa_sm100_fwd_decode_static(const fa_sm100_fwd_params* params, void* stream);
int64_t fa_sm100_static_scratch_bytes(void);

/* ---- Paged-KV decode (seqlen_q == 1) -----------------------------------------------------------------
 * SURVEY §8(f) rank 4; "paged KV cache" is a TODO of the reference (README.md:21), so the contract
 * is the contiguous decode above applied to the keys a block table names: sequence b owns
 * seqlens_k[b] keys, key j of it lives in physical page block_table[b * block_table_stride +
 * j / page_size] at in-page position j % page_size.  k_cache / v_cache element address of
 * (page, kv-head, in-page token, d) = page * kv_strides[0] + head * kv_strides[1] +
 * token * kv_strides[2] + d, which covers both [pages, page_size, Hkv, D] and
 * [pages, Hkv, page_size, D].  page_size must be a multiple of 16.  seqlens_k and block_table are
 * DEVICE int32 arrays (read by the kernel, no host sync); max_seqlen_k must be >= every
 * seqlens_k[b].  A sequence with no keys yields zeros.  Same numerics and workspace rules as
 * fa_sm100_fwd_decode.
 *
 * How the work is cut (the lengths are device data, the host cannot size a split per sequence):
 *   default — the CTAs of one wave (2 per SM) share the FLATTENED key space: every (sequence,
 *     kv-head) unit owns ceil(len / 64) blocks of 64 keys laid end to end and each CTA streams an
 *     equal run of blocks, whichever units they belong to; a unit cut between CTAs is merged by the
 *     last of them to arrive.  Every CTA reads the same number of bytes however ragged the batch
 *     is (one 16 K-key sequence among 31 of 512 keys: 4.2x faster than the split below).  Up to
 *     2048 sequences; the plan reports nsplit = 0 and chunk_keys = the smallest run of a CTA.
 *   FA_SM100_FLAG_PAGED_ONE_WAVE (or more than 2048 sequences) — every unit is cut into the same
 *     number of chunks, sized from max_seqlen_k so that one wave covers the GPU, exactly as
 *     fa_sm100_fwd_decode does.  3-10 % faster when all lengths are close to max_seqlen_k (chunk
 *     boundaries coincide with unit boundaries: fewer partial results), much slower when a few
 *     sequences are far longer than the rest (their CTAs finish last, alone).  A caller that has
 *     the lengths on the host can choose per batch: flash_attention_cute_b200/paged.py does. */
typedef struct fa_sm100_paged_decode_params {
    const void* q;        /* [B, Hq, D] element strides q_strides = {batch, head} */
    const void* k_cache;
    const void* v_cache;
    void* o;              /* [B, Hq, D] element strides o_strides = {batch, head} */
    const int32_t* block_table; /* device [B, block_table_stride] */
    const int32_t* seqlens_k;   /* device [B] */
    int64_t batch;
    int64_t heads_q;
    int64_t heads_kv;
    int64_t headdim;
    int64_t page_size;
    int64_t max_seqlen_k;
    int64_t block_table_stride;
    int64_t q_strides[2];
    int64_t o_strides[2];
    int64_t k_strides[3]; /* page, head, in-page token */
    int64_t v_strides[3];
    float softmax_scale;
    int32_t dtype;
    int32_t device;
    uint32_t flags;
} fa_sm100_paged_decode_params;

int fa_sm100_paged_decode_plan(const fa_sm100_paged_decode_params* params, int32_t* nsplit,
                               int32_t* chunk_keys, int64_t* workspace_bytes);
int fa_sm100_fwd_decode_paged(const fa_sm100_paged_decode_params* params, void* workspace,
                              int64_t workspace_bytes, void* stream);

/* ---- Variable-length (packed) batches ---------------------------------------------------------------
 * SURVEY §8(f) rank 4; the reference lists "varlen" as a TODO (README.md:18) and has no code for it,
 * so the semantics follow its fixed-length path sequence by sequence: sequence i owns the token rows
 * cu_seqlens_q[i] .. cu_seqlens_q[i+1] of q / o ([total_q, Hq, D]) and cu_seqlens_kv[i] ..
 * cu_seqlens_kv[i+1] of k / v ([total_kv, Hkv, D]) and is computed exactly like one batch entry of
 * fa_sm100_fwd (same GQA rule, same bottom-right causal alignment per sequence; bit-equal results).
 * cu_seqlens_* are DEVICE arrays of num_seqs + 1 non-decreasing int32 offsets starting at 0 and
 * ending at total_q / total_kv: nothing is read on the host, nothing synchronises, the call is
 * CUDA-graph capturable.  ONE persistent launch serves the whole batch (plus a one-CTA plan kernel
 * in front of it that orders the sequences longest first): the tile scheduler walks the offsets,
 * the tensor maps span the packed token axis, O rows are clipped per sequence, and V rows that
 * belong to a neighbouring sequence are zeroed in shared memory before they reach the tensor
 * core, so no value of another sequence can leak into a result.  Sequences with no queries are
 * skipped; a sequence with queries and no keys writes zeros.  Token strides must be multiples of
 * 8 elements.  max_seqlen_q is a hint (0 = unknown) that only steers the GQA packing choice.
 * The caller owns `workspace`: fa_sm100_varlen_workspace_bytes(num_seqs) bytes of device memory,
 * 16-byte aligned, contents irrelevant. */
typedef struct fa_sm100_varlen_params {
    const void* q; /* [total_q,  Hq,  D] */
    const void* k; /* [total_kv, Hkv, D] */
    const void* v;
    void* o;       /* [total_q,  Hq,  D] */
    const int32_t* cu_seqlens_q;  /* device, num_seqs + 1 entries */
    const int32_t* cu_seqlens_kv; /* device, num_seqs + 1 entries */
    int64_t num_seqs;
    int64_t total_q;
    int64_t total_kv;
    int64_t max_seqlen_q;
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

int64_t fa_sm100_varlen_workspace_bytes(int64_t num_seqs);
int fa_sm100_fwd_varlen(const fa_sm100_varlen_params* params, void* workspace, int64_t workspace_bytes,
                        void* stream);

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
 * call hit its in-kernel watchdog; clears the condition.  While the condition is set every entry
 * point above returns FA_SM100_ERR_KERNEL_ABORT without launching (the flag is mirrored into
 * host-mapped memory, so that check costs one host load), and launches already queued behind the
 * aborted one exit at once without touching their outputs. */
int fa_sm100_check_device(int device);

/* In-kernel watchdog: a wait inside a kernel gives up after `milliseconds` (default 2000; the
 * environment variable FA_SM100_WATCHDOG_MS sets the initial value; 0 disables the watchdog, e.g.
 * under a debugger or time-sliced MPS).  Process-wide, takes effect for later launches. */
void fa_sm100_set_watchdog_ms(int64_t milliseconds);

/* Library / kernel introspection used by tests and bench: number of kernel launches issued by
 * this library in this process, and a static description of the build. */
uint64_t fa_sm100_launch_count(void);
const char* fa_sm100_build_info(void);

/* Describe the variant fa_sm100_fwd would launch for `params` (no launch): fills grid size,
 * dynamic shared memory bytes, pack factor (1 = unpacked) and q rows per head per tile. */
int fa_sm100_plan(const fa_sm100_fwd_params* params, int32_t* grid, int32_t* smem_bytes,
                  int32_t* pack_g, int32_t* rows_per_head);
/* Name of the kernel variant fa_sm100_fwd launches for `params` ("" if they are invalid). */
const char* fa_sm100_kernel_name(const fa_sm100_fwd_params* params);

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
