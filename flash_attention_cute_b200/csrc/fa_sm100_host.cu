// Host side of the C ABI declared in include/fa_sm100.h: argument validation (the checks of
// reference csrc/flash_attention_api.cpp:17-59), variant selection (replaces
// csrc/kernel_dispatcher.h:20-52 and csrc/flash_attention_impl.cu:7-49), TMA tensor-map
// encoding, launch (replaces csrc/flash_attention_template.cuh:567-583).  No torch, no CUTLASS.
#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <mutex>

#include "../../include/fa_sm100.h"
#include "fa_sm100_kernel.cuh"
#include "fa_sm100_fwd.cuh"
#include "fa_sm100_fwd4.cuh"
#include "fa_sm100_fwd5.cuh"
#include "fa_sm100_fwd6.cuh"
#include "fa_sm100_decode.cuh"
#include "fa_sm100_rope.cuh"

namespace {

thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};
std::atomic<uint64_t> g_sched_seq{0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

// ---- driver entry point for cuTensorMapEncodeTiled (no link-time libcuda dependency) ----
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e =
            cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess) p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// ---- per-device state, created once ----
constexpr int kMaxDevices = 64;
constexpr int kSchedSlots = 256;  // launches that may be in flight at once on one device
struct DeviceState {
    std::once_flag once;
    int error = 0;
    int cc_major = 0, cc_minor = 0, num_sms = 0, max_smem_optin = 0;
    unsigned int* status = nullptr;  // 4 words of device memory, zeroed
    int* sched = nullptr;            // kSchedSlots x {next-tile counter, finished-CTA counter}, zeroed;
                                     // each launch takes the next slot, the kernel re-zeroes it on exit
    char msg[256] = "";
};
DeviceState g_dev[kMaxDevices];

DeviceState* device_state(int device) {
    if (device < 0 || device >= kMaxDevices) return nullptr;
    DeviceState* st = &g_dev[device];
    std::call_once(st->once, [st, device] {
        cudaError_t e;
        int prev = 0;
        cudaGetDevice(&prev);
        auto check = [&](cudaError_t err, const char* what) {
            if (err != cudaSuccess && st->error == 0) {
                st->error = FA_SM100_ERR_CUDA;
                snprintf(st->msg, sizeof(st->msg), "%s: %s", what, cudaGetErrorString(err));
            }
        };
        e = cudaSetDevice(device);
        check(e, "cudaSetDevice");
        if (e == cudaSuccess) {
            check(cudaDeviceGetAttribute(&st->cc_major, cudaDevAttrComputeCapabilityMajor, device), "attr");
            check(cudaDeviceGetAttribute(&st->cc_minor, cudaDevAttrComputeCapabilityMinor, device), "attr");
            check(cudaDeviceGetAttribute(&st->num_sms, cudaDevAttrMultiProcessorCount, device), "attr");
            check(cudaDeviceGetAttribute(&st->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device), "attr");
            check(cudaMalloc(&st->status, 4 * sizeof(unsigned int)), "cudaMalloc(status)");
            if (st->status) check(cudaMemset(st->status, 0, 4 * sizeof(unsigned int)), "cudaMemset(status)");
            check(cudaMalloc(&st->sched, kSchedSlots * 2 * sizeof(int)), "cudaMalloc(sched)");
            if (st->sched) check(cudaMemset(st->sched, 0, kSchedSlots * 2 * sizeof(int)), "cudaMemset(sched)");
        }
        cudaSetDevice(prev);
    });
    return st;
}

// ---- variant table ----
using KernelFn = void (*)(const fa100::KernelParams);
struct Variant {
    KernelFn fn;
    int smem_bytes;
    std::once_flag attr_once[kMaxDevices];
};

template <int kD, bool kBf16, int kStages>
Variant* variant() {
    static Variant v{fa100::fa_fwd_sm100_kernel<kD, kBf16, kStages>,
                     fa100::SmemLayout<kD, kStages>::kDynamicBytes};
    return &v;
}

constexpr int kStages128 = 5;  // 2 x 32 KB Q + 5 x 32 KB K/V = 224 KB
constexpr int kStages64 = 8;   // 2 x 16 KB Q + 8 x 16 KB K/V = 160 KB

Variant* pick_variant(int kd, bool bf16) {
    if (kd == 128) return bf16 ? variant<128, true, kStages128>() : variant<128, false, kStages128>();
    return bf16 ? variant<64, true, kStages64>() : variant<64, false, kStages64>();
}

// ---- generation-2 kernel variants (fa_sm100_fwd.cuh) ----
using KernelFn2 = void (*)(const fa100::FwdParams);
struct Variant2 {
    KernelFn2 fn;
    int smem_bytes;
    int threads;
    std::once_flag attr_once[kMaxDevices];
};
template <int kD, bool kBf16, int kStages, int kEmu, int kSplit, int kHalves, bool kTrace>
Variant2* variant2() {
    static Variant2 v{fa100::v2::fa_fwd_sm100_v2<kD, kBf16, kStages, kEmu, kSplit, kHalves, kTrace>,
                      fa100::v2::Smem<kD, kStages, kHalves>::kDynamicBytes,
                      fa100::v2::Cfg<kHalves>::kNumThreads};
    return &v;
}
// K/V ring depth: D=128 tiles are 32 KB (Q0+Q1 = 64 KB): 5 stages fit beside the 1-thread-per-row
// softmax state, 4 beside the row-max exchange buffers of the 2-threads-per-row layout.
template <int kD, int kHalves> constexpr int stages_for() { return kD == 64 ? 8 : (kHalves == 1 ? 5 : 4); }

// Defaults (chosen from the sweeps recorded in profiles/): threads per softmax row, emulated exp2
// pairs out of 4, split keys.
constexpr int kDefHalves = 2, kDefEmu = 0, kDefSplit = 128;

template <int kD, bool kBf16>
Variant2* pick_tuned(int halves, int emu, int split, bool trace) {
#ifdef FA_SM100_TUNE
#define FA_V(H, E, S)                                                                            \
    if (halves == H && emu == E && split == S)                                                   \
        return trace ? variant2<kD, kBf16, stages_for<kD, H>(), E, S, H, true>()                 \
                     : variant2<kD, kBf16, stages_for<kD, H>(), E, S, H, false>();
    FA_V(1, 0, 128) FA_V(2, 0, 128)
#undef FA_V
#endif
    (void)halves; (void)emu; (void)split; (void)trace;
    return nullptr;
}

// ---- generation-4 kernel variants (fa_sm100_fwd4.cuh) ----
template <int kD, bool kBf16, int kEmu, int kSplit, bool kTrace>
Variant2* variant4() {
    constexpr int kStages = kD == 64 ? 8 : 5;
    static Variant2 v{fa100::v4::fa_fwd_sm100_v4<kD, kBf16, kStages, kEmu, kSplit, kTrace>,
                      fa100::v4::Smem<kD, kStages>::kDynamicBytes, fa100::v4::Cfg::kNumThreads};
    return &v;
}
constexpr int kDefEmu4 = 0, kDefSplit4 = 128;
Variant2* pick_variant4(int kd, bool bf16, int emu, int split, bool trace) {
    if (emu < 0) emu = kDefEmu4;
    if (split < 0) split = kDefSplit4;
#ifdef FA_SM100_TUNE
#define FA_V4(D, BF, E, S)                                                                 \
    if (kd == D && bf16 == BF && emu == E && split == S)                                   \
        return trace ? variant4<D, BF, E, S, true>() : variant4<D, BF, E, S, false>();
    FA_V4(128, true, 0, 128) FA_V4(128, true, 1, 128) FA_V4(128, true, 2, 128) FA_V4(128, true, 3, 128)
    FA_V4(128, true, 0, 96) FA_V4(128, true, 1, 96) FA_V4(128, true, 2, 96) FA_V4(128, true, 3, 96)
    FA_V4(128, true, 0, 64) FA_V4(128, true, 2, 64)
#undef FA_V4
#endif
    (void)trace;
    constexpr int E = kDefEmu4, S = kDefSplit4;
    if (kd == 128) return bf16 ? variant4<128, true, E, S, false>() : variant4<128, false, E, S, false>();
    return bf16 ? variant4<64, true, E, S, false>() : variant4<64, false, E, S, false>();
}

// ---- generation-5 kernel variants (fa_sm100_fwd5.cuh): half-block pipeline ----
template <int kD, bool kBf16, int kEmu, bool kSpec, bool kTrace>
Variant2* variant5() {
    constexpr int kStages = kD == 64 ? 8 : 5;  // ring stages of two 64-key half tiles of K / V
    static Variant2 v{fa100::v5::fa_fwd_sm100_v5<kD, kBf16, kStages, kEmu, kSpec, kTrace>,
                      fa100::v5::Smem<kD, kStages>::kDynamicBytes, fa100::v5::Cfg::kNumThreads};
    return &v;
}
// Defaults from the interleaved A/B sweep in profiles/r01/ab_v5_emu_spec.log: D=128 -> 3 of 8 pairs on
// the FMA pipe, plain two-pass softmax; D=64 -> 2 of 8, speculative stale-max softmax.
constexpr int kDefEmu5_128 = 3, kDefEmu5_64 = 2;
constexpr bool kDefSpec5_128 = false, kDefSpec5_64 = true;
// split flag 64 selects the speculative-max softmax, 128 the two-pass one (tuning builds)
Variant2* pick_variant5(int kd, bool bf16, int emu, int split, bool trace) {
    if (emu < 0) emu = kd == 128 ? kDefEmu5_128 : kDefEmu5_64;
    const bool spec = split < 0 ? (kd == 128 ? kDefSpec5_128 : kDefSpec5_64) : split == 64;
#ifdef FA_SM100_TUNE
#define FA_V5(D, BF, E, SP)                                                                \
    if (kd == D && bf16 == BF && emu == E && spec == SP)                                   \
        return trace ? variant5<D, BF, E, SP, true>() : variant5<D, BF, E, SP, false>();
    FA_V5(128, true, 0, false) FA_V5(128, true, 1, false) FA_V5(128, true, 2, false) FA_V5(128, true, 3, false)
    FA_V5(128, true, 1, true) FA_V5(128, true, 2, true) FA_V5(128, true, 3, true)
    FA_V5(64, false, 1, false) FA_V5(64, false, 2, false) FA_V5(64, false, 3, false)
    FA_V5(64, false, 2, true) FA_V5(64, false, 3, true)
#undef FA_V5
#endif
    (void)trace;
    if (kd == 128)
        return bf16 ? variant5<128, true, kDefEmu5_128, kDefSpec5_128, false>()
                    : variant5<128, false, kDefEmu5_128, kDefSpec5_128, false>();
    return bf16 ? variant5<64, true, kDefEmu5_64, kDefSpec5_64, false>()
                : variant5<64, false, kDefEmu5_64, kDefSpec5_64, false>();
}

// ---- generation-6 kernel variants (fa_sm100_fwd6.cuh): 128-key blocks, P handed over in parts ----
template <int kD, bool kBf16, int kEmu, int kParts, bool kTrace>
Variant2* variant6() {
    constexpr int kStages = kD == 64 ? 8 : 5;
    static Variant2 v{fa100::v6::fa_fwd_sm100_v6<kD, kBf16, kStages, kEmu, kParts, kTrace>,
                      fa100::v6::Smem<kD, kStages>::kDynamicBytes, fa100::v6::Cfg::kNumThreads};
    return &v;
}
constexpr int kDefEmu6 = 2, kDefParts6 = 2;
// split flag 64 -> 4 parts, 96 -> 2 parts, 128 -> 1 part
Variant2* pick_variant6(int kd, bool bf16, int emu, int split, bool trace) {
    if (emu < 0) emu = kDefEmu6;
    const int parts = split < 0 ? kDefParts6 : (split == 64 ? 4 : split == 96 ? 2 : 1);
#ifdef FA_SM100_TUNE
#define FA_V6(D, BF, E, PT)                                                                \
    if (kd == D && bf16 == BF && emu == E && parts == PT)                                  \
        return trace ? variant6<D, BF, E, PT, true>() : variant6<D, BF, E, PT, false>();
    FA_V6(128, true, 0, 1) FA_V6(128, true, 0, 2) FA_V6(128, true, 0, 4)
    FA_V6(128, true, 2, 1) FA_V6(128, true, 2, 2) FA_V6(128, true, 2, 4)
    FA_V6(128, true, 3, 2) FA_V6(128, true, 3, 4) FA_V6(128, true, 1, 2) FA_V6(128, true, 1, 4)
#undef FA_V6
#endif
    (void)trace;
    constexpr int E = kDefEmu6, PT = kDefParts6;
    if (kd == 128) return bf16 ? variant6<128, true, E, PT, false>() : variant6<128, false, E, PT, false>();
    return bf16 ? variant6<64, true, E, PT, false>() : variant6<64, false, E, PT, false>();
}

Variant2* pick_variant2(int kd, bool bf16, int halves, int emu, int split, bool trace) {
    if (halves < 0) halves = kDefHalves;
    if (emu < 0) emu = kDefEmu;
    if (split < 0 || halves == 2) split = halves == 2 ? 128 : (split < 0 ? kDefSplit : split);
    Variant2* v = nullptr;
    if (kd == 128 && bf16) v = pick_tuned<128, true>(halves, emu, split, trace);
    if (kd == 64 && !bf16) v = pick_tuned<64, false>(halves, emu, split, trace);
    if (v) return v;
    constexpr int H = kDefHalves, E = kDefEmu, S = kDefSplit;
    if (kd == 128)
        return bf16 ? variant2<128, true, stages_for<128, H>(), E, S, H, false>()
                    : variant2<128, false, stages_for<128, H>(), E, S, H, false>();
    return bf16 ? variant2<64, true, stages_for<64, H>(), E, S, H, false>()
                : variant2<64, false, stages_for<64, H>(), E, S, H, false>();
}

constexpr int kDefaultGenStatic = 5;  // true: generation 4 (fa_sm100_fwd4.cuh) is what fa_sm100_fwd launches
// FA_SM100_GEN=2|4 in the environment overrides the default generation (test / tuning aid).
const int kDefaultGen = [] {
    const char* e = getenv("FA_SM100_GEN");
    if (e && (e[0] == '2' || e[0] == '4' || e[0] == '5' || e[0] == '6')) return e[0] - '0';
    return kDefaultGenStatic;
}();
long long* g_trace = nullptr;  // debug timeline buffer (fa_sm100_internal_set_trace)

struct Plan {
    int kd;             // kernel head-dim variant (64 or 128)
    int pack_g;         // 1 = unpacked
    int rows_per_head;  // P
    int num_m_blocks;
    int head_units;
    int grid;
    Variant* var;
};

int validate_and_plan(const fa_sm100_fwd_params* p, Plan* plan) {
    if (!p) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "params is NULL");
    if (!p->q || !p->k || !p->v || !p->o)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "q, k, v, o must be non-NULL device pointers");
    // messages mirror reference csrc/flash_attention_api.cpp:22-55
    if (p->batch <= 0 || p->heads_q <= 0 || p->seqlen_q <= 0 || p->headdim <= 0 ||
        p->heads_kv <= 0 || p->seqlen_kv <= 0)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "q, k, v must have at least one element");
    if (p->heads_q < p->heads_kv)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT,
                    "number of heads in q must be greater or equal to number of heads in k and v");
    if (p->heads_q % p->heads_kv != 0)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT,
                    "number of heads in q must be multiple of number of heads in k and v");
    if (p->dtype != FA_SM100_FP16 && p->dtype != FA_SM100_BF16)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "q, k, v only support fp16 or bf16 data type");
    if (p->headdim % 8 != 0)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "hidden dimension must be multiple of 8");
    if (p->headdim > 128)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "only support hidden dimension <= 128");
    if (p->seqlen_q > (1ll << 30) || p->seqlen_kv > (1ll << 30) || p->batch > 65535 * 16 ||
        p->heads_q > 65535)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "problem size out of range");
    auto aligned16 = [](const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15u) == 0; };
    if (!aligned16(p->q) || !aligned16(p->k) || !aligned16(p->v) || !aligned16(p->o))
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "q, k, v, o must be 16-byte aligned");
    auto strides_ok = [](const int64_t* s, const int64_t* dims) {
        for (int i = 0; i < 3; ++i)
            if (dims[i] > 1 && (s[i] <= 0 || s[i] % 8 != 0)) return false;
        return true;
    };
    const int64_t qd[3] = {p->batch, p->heads_q, p->seqlen_q};
    const int64_t kd[3] = {p->batch, p->heads_kv, p->seqlen_kv};
    if (!strides_ok(p->q_strides, qd) || !strides_ok(p->k_strides, kd) ||
        !strides_ok(p->v_strides, kd) || !strides_ok(p->o_strides, qd))
        return fail(FA_SM100_ERR_INVALID_ARGUMENT,
                    "batch/head/sequence strides must be positive multiples of 8 elements (16 bytes)");

    const int group = int(p->heads_q / p->heads_kv);
    plan->kd = p->headdim <= 64 ? 64 : 128;
    bool pack = group > 1 && group <= 128;
    if (p->flags & FA_SM100_FLAG_NO_PACK_GQA) pack = false;
    // Heuristic: packing pays when one head alone cannot fill the CTA's 256 query rows; for long
    // prefill both mappings do the same MMA work and the unpacked one has the simpler mask.
    if (pack && !(p->flags & FA_SM100_FLAG_FORCE_PACK_GQA) && p->seqlen_q >= 256) pack = false;
    plan->pack_g = pack ? group : 1;
    plan->rows_per_head = pack ? 128 / group : 128;
    plan->num_m_blocks = int((p->seqlen_q + 2 * plan->rows_per_head - 1) / (2 * plan->rows_per_head));
    plan->head_units = int(pack ? p->heads_kv : p->heads_q);
    const int64_t grid = int64_t(plan->num_m_blocks) * plan->head_units * p->batch;
    if (grid > 0x7fffffffll) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "grid too large");
    plan->grid = int(grid);
    plan->var = pick_variant(plan->kd, p->dtype == FA_SM100_BF16);
    return FA_SM100_OK;
}

int encode_map(CUtensorMap* map, CUtensorMapDataType dt, int rank, const void* base,
               const cuuint64_t* dims, const cuuint64_t* strides_bytes, const cuuint32_t* box,
               const char* what) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return fail(FA_SM100_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = fn(map, dt, cuuint32_t(rank), const_cast<void*>(base), dims, strides_bytes, box,
                    estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(FA_SM100_ERR_CUDA, "cuTensorMapEncodeTiled(%s) failed with CUresult %d", what, int(r));
    return FA_SM100_OK;
}

// A size-1 dimension may carry any stride in torch; TMA still wants a 16 B multiple.
inline cuuint64_t stride_bytes(int64_t stride_elems, int64_t dim, cuuint64_t fallback) {
    if (dim <= 1 || stride_elems <= 0) return fallback;
    return cuuint64_t(stride_elems) * 2;
}

}  // namespace

extern "C" const char* fa_sm100_last_error(void) { return g_err; }

extern "C" uint64_t fa_sm100_launch_count(void) { return g_launches.load(); }

extern "C" const char* fa_sm100_build_info(void) {
    return "fa_sm100: sm_100a tcgen05/TMEM/TMA FlashAttention-2 forward; default kernel "
           "fa100::v5::fa_fwd_sm100_v5 (persistent, 512 threads/CTA, 2x128-row M tiles, 64-key half "
           "blocks, 2 S buffers per tile); variants D{64,128} x {fp16,bf16}";
}

extern "C" int fa_sm100_plan(const fa_sm100_fwd_params* params, int32_t* grid, int32_t* smem_bytes,
                             int32_t* pack_g, int32_t* rows_per_head) {
    Plan plan;
    int rc = validate_and_plan(params, &plan);
    if (rc) return rc;
    if (grid) *grid = plan.grid;
    if (smem_bytes) *smem_bytes = plan.var->smem_bytes;
    if (pack_g) *pack_g = plan.pack_g;
    if (rows_per_head) *rows_per_head = plan.rows_per_head;
    return FA_SM100_OK;
}

extern "C" int fa_sm100_fwd(const fa_sm100_fwd_params* p, void* stream) {
    Plan plan;
    int rc = validate_and_plan(p, &plan);
    if (rc) return rc;

    DeviceState* st = device_state(p->device);
    if (!st) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "bad device ordinal %d", p->device);
    if (st->error) return fail(st->error, "device init failed: %s", st->msg);
    if (st->cc_major != 10)
        return fail(FA_SM100_ERR_UNSUPPORTED_DEVICE,
                    "fa_sm100 needs an sm_100 (B200) GPU; device %d is sm_%d%d — there is no fallback",
                    p->device, st->cc_major, st->cc_minor);

    int prev = 0;
    cudaGetDevice(&prev);
    if (prev != p->device) cudaSetDevice(p->device);
    struct Restore {
        int prev, cur;
        ~Restore() { if (prev != cur) cudaSetDevice(prev); }
    } restore{prev, p->device};

    const bool use_v1 = (p->flags & FA_SM100_FLAG_KERNEL_V1) != 0;
    const int tune_emu = int((p->flags >> 8) & 0xfu) - 1;                 // -1 = default
    static const int kSplitSel[4] = {-1, 64, 96, 128};
    const int tune_split = kSplitSel[(p->flags >> 12) & 0x3u];
    Variant* var = plan.var;
    const int tune_halves = int((p->flags >> 14) & 0x3u) == 0 ? -1 : int((p->flags >> 14) & 0x3u);
    const bool use_v4 = !use_v1 && tune_halves == 3;
    const bool use_v5 = !use_v1 && ((tune_halves < 0 && kDefaultGen == 5) || tune_halves == 2);
    const bool use_v6 = !use_v1 && tune_halves < 0 && kDefaultGen == 6;
    Variant2* var2 = use_v1 ? nullptr
                   : use_v6 ? pick_variant6(plan.kd, p->dtype == FA_SM100_BF16, tune_emu, tune_split, g_trace != nullptr)
                   : use_v5 ? pick_variant5(plan.kd, p->dtype == FA_SM100_BF16, tune_emu, tune_split, g_trace != nullptr)
                   : use_v4 ? pick_variant4(plan.kd, p->dtype == FA_SM100_BF16, tune_emu, tune_split, g_trace != nullptr)
                            : pick_variant2(plan.kd, p->dtype == FA_SM100_BF16, tune_halves, tune_emu,
                                            tune_split, g_trace != nullptr);
    const int smem_bytes = use_v1 ? var->smem_bytes : var2->smem_bytes;
    if (smem_bytes > st->max_smem_optin)
        return fail(FA_SM100_ERR_UNSUPPORTED_DEVICE, "kernel needs %d B shared memory, device allows %d",
                    smem_bytes, st->max_smem_optin);
    cudaError_t attr_err = cudaSuccess;
    if (use_v1) {
        std::call_once(var->attr_once[p->device], [&] {
            attr_err = cudaFuncSetAttribute(reinterpret_cast<const void*>(var->fn),
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, var->smem_bytes);
        });
    } else {
        std::call_once(var2->attr_once[p->device], [&] {
            attr_err = cudaFuncSetAttribute(reinterpret_cast<const void*>(var2->fn),
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, var2->smem_bytes);
        });
    }
    if (attr_err != cudaSuccess)
        return fail(FA_SM100_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));

    fa100::FwdParams kp;
    memset(&kp, 0, sizeof(kp));
    const CUtensorMapDataType dt =
        p->dtype == FA_SM100_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    const int group = int(p->heads_q / p->heads_kv);
    const cuuint64_t D = cuuint64_t(p->headdim);
    {
        // Q as (d, s, g, hkv, b): head h = hkv * G + g
        const cuuint64_t row = D * 2;
        cuuint64_t s_s = stride_bytes(p->q_strides[2], p->seqlen_q, row);
        cuuint64_t s_g = stride_bytes(p->q_strides[1], p->heads_q, s_s * cuuint64_t(p->seqlen_q));
        cuuint64_t s_h = s_g * cuuint64_t(group);
        cuuint64_t s_b = stride_bytes(p->q_strides[0], p->batch, s_h * cuuint64_t(p->heads_kv));
        cuuint64_t dims[5] = {D, cuuint64_t(p->seqlen_q), cuuint64_t(group), cuuint64_t(p->heads_kv),
                              cuuint64_t(p->batch)};
        cuuint64_t strides[4] = {s_s, s_g, s_h, s_b};
        cuuint32_t box[5] = {64, cuuint32_t(plan.rows_per_head), cuuint32_t(plan.pack_g), 1, 1};
        rc = encode_map(&kp.tm_q, dt, 5, p->q, dims, strides, box, "q");
        if (rc) return rc;
    }
    auto encode_kv = [&](CUtensorMap* m, const void* base, const int64_t* es, const char* what) {
        const cuuint64_t row = D * 2;
        cuuint64_t s_s = stride_bytes(es[2], p->seqlen_kv, row);
        cuuint64_t s_h = stride_bytes(es[1], p->heads_kv, s_s * cuuint64_t(p->seqlen_kv));
        cuuint64_t s_b = stride_bytes(es[0], p->batch, s_h * cuuint64_t(p->heads_kv));
        cuuint64_t dims[4] = {D, cuuint64_t(p->seqlen_kv), cuuint64_t(p->heads_kv), cuuint64_t(p->batch)};
        cuuint64_t strides[3] = {s_s, s_h, s_b};
        cuuint32_t box[4] = {64, cuuint32_t(use_v5 ? 64 : 128), 1, 1};
        return encode_map(m, dt, 4, base, dims, strides, box, what);
    };
    rc = encode_kv(&kp.tm_k, p->k, p->k_strides, "k");
    if (rc) return rc;
    rc = encode_kv(&kp.tm_v, p->v, p->v_strides, "v");
    if (rc) return rc;

    kp.o = p->o;
    kp.o_stride_b = p->o_strides[0];
    kp.o_stride_h = p->o_strides[1];
    kp.o_stride_s = p->o_strides[2];
    kp.batch = int(p->batch);
    kp.heads_q = int(p->heads_q);
    kp.heads_kv = int(p->heads_kv);
    kp.group = group;
    kp.seqlen_q = int(p->seqlen_q);
    kp.seqlen_k = int(p->seqlen_kv);
    kp.headdim = int(p->headdim);
    kp.pack_g = plan.pack_g;
    kp.rows_per_head = plan.rows_per_head;
    kp.num_m_blocks = plan.num_m_blocks;
    kp.head_units = plan.head_units;
    kp.total_heads = plan.head_units * int(p->batch);
    kp.num_tiles = kp.num_m_blocks * kp.total_heads;
    kp.q_tile_bytes = (plan.kd / 64) * 64 * plan.rows_per_head * plan.pack_g * 2;
    kp.causal = p->causal ? 1 : 0;
    kp.scale_log2 = p->softmax_scale * float(M_LOG2E);
    kp.status = st->status;
    kp.trace = g_trace;
    kp.sched = st->sched + 2 * (g_sched_seq.fetch_add(1) % kSchedSlots);
    {
        // L2 sections: as many consecutive (batch, head) units as keep their K+V under ~40 MB
        // (a third of the 126 MB L2), in whole kv groups when the q heads are not packed.
        const double kv_bytes_per_kvhead = 2.0 * double(p->seqlen_kv) * double(plan.kd) * 2.0;
        const int units_per_kvhead = plan.pack_g > 1 ? 1 : group;
        int kvheads = int(40.0 * 1024 * 1024 / kv_bytes_per_kvhead);
        if (kvheads < 1) kvheads = 1;
        long long sec = (long long)kvheads * units_per_kvhead;
        if (sec > kp.total_heads) sec = kp.total_heads;
        kp.section_heads = int(sec);
    }

    if (use_v1) {
        fa100::KernelParams k1;
        memset(&k1, 0, sizeof(k1));
        k1.tm_q = kp.tm_q; k1.tm_k = kp.tm_k; k1.tm_v = kp.tm_v;
        k1.o = kp.o; k1.o_stride_b = kp.o_stride_b; k1.o_stride_h = kp.o_stride_h; k1.o_stride_s = kp.o_stride_s;
        k1.batch = kp.batch; k1.heads_q = kp.heads_q; k1.heads_kv = kp.heads_kv; k1.group = kp.group;
        k1.seqlen_q = kp.seqlen_q; k1.seqlen_k = kp.seqlen_k; k1.headdim = kp.headdim;
        k1.pack_g = kp.pack_g; k1.rows_per_head = kp.rows_per_head; k1.num_m_blocks = kp.num_m_blocks;
        k1.head_units = kp.head_units; k1.q_tile_bytes = kp.q_tile_bytes; k1.causal = kp.causal;
        k1.scale_log2 = kp.scale_log2; k1.status = kp.status;
        var->fn<<<dim3(plan.grid), dim3(fa100::kNumThreads), var->smem_bytes,
                  static_cast<cudaStream_t>(stream)>>>(k1);
    } else {
        const int grid = kp.num_tiles < st->num_sms ? kp.num_tiles : st->num_sms;
        var2->fn<<<dim3(grid), dim3(var2->threads), var2->smem_bytes,
                   static_cast<cudaStream_t>(stream)>>>(kp);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(FA_SM100_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
    g_launches.fetch_add(1);
    return FA_SM100_OK;
}

// ------------------------------------------------------------------------------------------------
// Split-KV decode (Sq == 1): fa_sm100_decode.cuh
// ------------------------------------------------------------------------------------------------
namespace {

struct DecodePlan {
    int kd, group, units, nsplit, chunk;
    int64_t counters_bytes, ml_bytes, o_bytes;
};

int decode_plan(const fa_sm100_fwd_params* p, int num_sms, DecodePlan* dp) {
    Plan plan;
    int rc = validate_and_plan(p, &plan);
    if (rc) return rc;
    if (p->seqlen_q != 1)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "decode kernel needs seqlen_q == 1 (got %lld)",
                    (long long)p->seqlen_q);
    dp->group = int(p->heads_q / p->heads_kv);
    if (dp->group > fa100::decode::kRows)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT,
                    "decode kernel packs at most %d q heads per kv head (got %d); use fa_sm100_fwd",
                    fa100::decode::kRows, dp->group);
    if (!(p->softmax_scale > 0.f))
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "decode kernel needs softmax_scale > 0");
    if (p->seqlen_kv > 0x7fffffffll / 2 || p->batch * p->heads_kv > (1ll << 24))
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "problem size out of range");
    dp->kd = plan.kd;
    dp->units = int(p->batch * p->heads_kv);
    // One wave: two CTAs per SM.  A split shorter than 256 keys is not worth its merge.
    const int tile = fa100::decode::kWarps * fa100::decode::kTileKeys;
    const int slots = 2 * (num_sms > 0 ? num_sms : 148);
    int64_t nsplit = slots / dp->units;
    const int64_t max_split = p->seqlen_kv / 256;
    if (nsplit > max_split) nsplit = max_split;
    if (nsplit < 1) nsplit = 1;
    int64_t chunk = (p->seqlen_kv + nsplit - 1) / nsplit;
    chunk = (chunk + tile - 1) / tile * tile;
    nsplit = (p->seqlen_kv + chunk - 1) / chunk;
    dp->nsplit = int(nsplit);
    dp->chunk = int(chunk);
    dp->counters_bytes = nsplit > 1 ? (int64_t(dp->units) * 4 + 255) / 256 * 256 : 0;
    dp->ml_bytes = nsplit > 1 ? int64_t(dp->units) * nsplit * fa100::decode::kRows * 2 * 4 : 0;
    dp->o_bytes = nsplit > 1 ? int64_t(dp->units) * nsplit * fa100::decode::kRows * dp->kd * 4 : 0;
    return FA_SM100_OK;
}

using DecodeFn = void (*)(const fa100::decode::DecodeParams);
struct DecodeVariant {
    DecodeFn fn;
    int smem_bytes;
    std::once_flag attr_once[kMaxDevices];
};
template <int kD, bool kBf16>
DecodeVariant* decode_variant() {
    static DecodeVariant v{fa100::decode::fa_decode_sm100<kD, kBf16>, fa100::decode::Smem<kD>::kTotal};
    return &v;
}

}  // namespace

extern "C" int fa_sm100_decode_plan(const fa_sm100_fwd_params* p, int32_t* nsplit, int32_t* chunk,
                                    int64_t* workspace_bytes) {
    DecodePlan dp;
    int num_sms = 148;
    if (p) {
        DeviceState* st = (p->device >= 0 && p->device < kMaxDevices) ? &g_dev[p->device] : nullptr;
        if (st && st->num_sms > 0) num_sms = st->num_sms;  // only if the device was initialised already
    }
    int rc = decode_plan(p, num_sms, &dp);
    if (rc) return rc;
    if (nsplit) *nsplit = dp.nsplit;
    if (chunk) *chunk = dp.chunk;
    if (workspace_bytes) *workspace_bytes = dp.counters_bytes + dp.ml_bytes + dp.o_bytes;
    return FA_SM100_OK;
}

extern "C" int fa_sm100_fwd_decode(const fa_sm100_fwd_params* p, void* workspace,
                                   int64_t workspace_bytes, void* stream) {
    if (!p) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "params is NULL");
    DeviceState* st = device_state(p->device);
    if (!st) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "bad device ordinal %d", p->device);
    if (st->error) return fail(st->error, "device init failed: %s", st->msg);
    if (st->cc_major != 10)
        return fail(FA_SM100_ERR_UNSUPPORTED_DEVICE,
                    "fa_sm100 needs an sm_100 (B200) GPU; device %d is sm_%d%d — there is no fallback",
                    p->device, st->cc_major, st->cc_minor);
    DecodePlan dp;
    int rc = decode_plan(p, st->num_sms, &dp);
    if (rc) return rc;
    const int64_t need = dp.counters_bytes + dp.ml_bytes + dp.o_bytes;
    if (dp.nsplit > 1) {
        if (!workspace || workspace_bytes < need)
            return fail(FA_SM100_ERR_INVALID_ARGUMENT,
                        "decode workspace too small: need %lld bytes (fa_sm100_decode_plan), got %lld",
                        (long long)need, (long long)workspace_bytes);
        if (reinterpret_cast<uintptr_t>(workspace) & 255u)
            return fail(FA_SM100_ERR_INVALID_ARGUMENT, "decode workspace must be 256-byte aligned");
    }

    int prev = 0;
    cudaGetDevice(&prev);
    if (prev != p->device) cudaSetDevice(p->device);
    struct Restore {
        int prev, cur;
        ~Restore() { if (prev != cur) cudaSetDevice(prev); }
    } restore{prev, p->device};

    const bool bf16 = p->dtype == FA_SM100_BF16;
    DecodeVariant* var = dp.kd == 128 ? (bf16 ? decode_variant<128, true>() : decode_variant<128, false>())
                                      : (bf16 ? decode_variant<64, true>() : decode_variant<64, false>());
    cudaError_t attr_err = cudaSuccess;
    std::call_once(var->attr_once[p->device], [&] {
        attr_err = cudaFuncSetAttribute(reinterpret_cast<const void*>(var->fn),
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, var->smem_bytes);
    });
    if (attr_err != cudaSuccess)
        return fail(FA_SM100_ERR_CUDA, "cudaFuncSetAttribute(decode): %s", cudaGetErrorString(attr_err));

    fa100::decode::DecodeParams kp;
    memset(&kp, 0, sizeof(kp));
    kp.q = p->q; kp.k = p->k; kp.v = p->v; kp.o = p->o;
    kp.batch = int(p->batch); kp.heads_kv = int(p->heads_kv); kp.group = dp.group;
    kp.seqlen_k = int(p->seqlen_kv); kp.headdim = int(p->headdim);
    kp.q_sb = p->q_strides[0]; kp.q_sh = p->q_strides[1];
    kp.k_sb = p->k_strides[0]; kp.k_sh = p->k_strides[1]; kp.k_ss = p->k_strides[2];
    kp.v_sb = p->v_strides[0]; kp.v_sh = p->v_strides[1]; kp.v_ss = p->v_strides[2];
    kp.o_sb = p->o_strides[0]; kp.o_sh = p->o_strides[1];
    kp.scale_log2 = p->softmax_scale * 1.4426950408889634f;
    kp.nsplit = dp.nsplit; kp.chunk = dp.chunk;
    if (dp.nsplit > 1) {
        uint8_t* ws = static_cast<uint8_t*>(workspace);
        kp.counters = reinterpret_cast<int*>(ws);
        kp.part_ml = reinterpret_cast<float*>(ws + dp.counters_bytes);
        kp.part_o = reinterpret_cast<float*>(ws + dp.counters_bytes + dp.ml_bytes);
    }
    var->fn<<<dim3(dp.units * dp.nsplit), dim3(fa100::decode::kThreads), var->smem_bytes,
              static_cast<cudaStream_t>(stream)>>>(kp);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(FA_SM100_ERR_CUDA, "decode kernel launch: %s", cudaGetErrorString(e));
    g_launches.fetch_add(1);
    return FA_SM100_OK;
}

// ---------------------------------------------------------------------------------------------
// Variable-length batches: one forward launch per sequence (include/fa_sm100.h)
// ---------------------------------------------------------------------------------------------
extern "C" int fa_sm100_fwd_varlen(const fa_sm100_varlen_params* p, void* stream) {
    if (!p) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "params is NULL");
    if (!p->cu_seqlens_q || !p->cu_seqlens_kv)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "cu_seqlens_q / cu_seqlens_kv must be host arrays");
    if (p->num_seqs < 0) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "num_seqs must be >= 0");
    if (p->cu_seqlens_q[0] != 0 || p->cu_seqlens_kv[0] != 0)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "cu_seqlens must start at 0");
    for (int64_t i = 0; i < p->num_seqs; ++i)
        if (p->cu_seqlens_q[i + 1] < p->cu_seqlens_q[i] || p->cu_seqlens_kv[i + 1] < p->cu_seqlens_kv[i])
            return fail(FA_SM100_ERR_INVALID_ARGUMENT, "cu_seqlens must be non-decreasing (sequence %lld)",
                        (long long)i);
    const int64_t ts[4] = {p->q_strides[0], p->k_strides[0], p->v_strides[0], p->o_strides[0]};
    for (int j = 0; j < 4; ++j)
        if (ts[j] % 8 != 0)
            return fail(FA_SM100_ERR_INVALID_ARGUMENT,
                        "varlen token strides must be multiples of 8 elements (16 bytes)");
    for (int64_t i = 0; i < p->num_seqs; ++i) {
        const int64_t q0 = p->cu_seqlens_q[i], k0 = p->cu_seqlens_kv[i];
        const int64_t sq = p->cu_seqlens_q[i + 1] - q0, sk = p->cu_seqlens_kv[i + 1] - k0;
        if (sq == 0) continue;
        if (sk == 0) {
            // no visible key: the fixed-length path yields zeros (sum == 0 -> scale 1 on a zero accumulator)
            cudaError_t e = cudaMemset2DAsync(
                static_cast<char*>(p->o) + q0 * p->o_strides[0] * 2, size_t(p->o_strides[0]) * 2, 0,
                size_t(p->heads_q * p->headdim) * 2, size_t(sq), static_cast<cudaStream_t>(stream));
            if (p->o_strides[1] != p->headdim || e != cudaSuccess)
                return fail(e != cudaSuccess ? FA_SM100_ERR_CUDA : FA_SM100_ERR_INVALID_ARGUMENT,
                            "sequence %lld has queries but no keys (needs o head stride == D to zero-fill)",
                            (long long)i);
            continue;
        }
        fa_sm100_fwd_params f;
        memset(&f, 0, sizeof(f));
        f.q = static_cast<const char*>(p->q) + q0 * p->q_strides[0] * 2;
        f.k = static_cast<const char*>(p->k) + k0 * p->k_strides[0] * 2;
        f.v = static_cast<const char*>(p->v) + k0 * p->v_strides[0] * 2;
        f.o = static_cast<char*>(p->o) + q0 * p->o_strides[0] * 2;
        f.batch = 1;
        f.heads_q = p->heads_q;
        f.heads_kv = p->heads_kv;
        f.seqlen_q = sq;
        f.seqlen_kv = sk;
        f.headdim = p->headdim;
        const int64_t* src[4] = {p->q_strides, p->k_strides, p->v_strides, p->o_strides};
        int64_t* dst[4] = {f.q_strides, f.k_strides, f.v_strides, f.o_strides};
        const int64_t rows[4] = {sq, sk, sk, sq};
        const int64_t heads[4] = {p->heads_q, p->heads_kv, p->heads_kv, p->heads_q};
        for (int j = 0; j < 4; ++j) {
            dst[j][2] = src[j][0];
            dst[j][1] = src[j][1];
            // batch == 1: any value works; keep it consistent with a dense tensor
            dst[j][0] = (rows[j] * src[j][0] > heads[j] * src[j][1]) ? rows[j] * src[j][0] : heads[j] * src[j][1];
        }
        f.softmax_scale = p->softmax_scale;
        f.causal = p->causal;
        f.dtype = p->dtype;
        f.device = p->device;
        f.flags = p->flags;
        int rc = fa_sm100_fwd(&f, stream);
        if (rc) return rc;
    }
    return FA_SM100_OK;
}

// ------------------------------------------------------------------------------------------------
// Fused RoPE on q and k: fa_sm100_rope.cuh
// ------------------------------------------------------------------------------------------------
extern "C" int fa_sm100_rope(const fa_sm100_rope_params* p, void* stream) {
    if (!p) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "params is NULL");
    if (!p->q || !p->k || !p->q_out || !p->k_out || !p->cos || !p->sin)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "q, k, q_out, k_out, cos, sin must be non-NULL device pointers");
    if (p->batch <= 0 || p->heads_q <= 0 || p->heads_kv <= 0 || p->seqlen <= 0 || p->headdim <= 0)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "q, k must have at least one element");
    if (p->headdim % 16 != 0 || p->headdim > 512)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "rope: hidden dimension must be a multiple of 16 and <= 512");
    if (p->dtype != FA_SM100_FP16 && p->dtype != FA_SM100_BF16)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "q, k only support fp16 or bf16 data type");
    auto aligned16 = [](const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15u) == 0; };
    if (!aligned16(p->q) || !aligned16(p->k) || !aligned16(p->q_out) || !aligned16(p->k_out) ||
        !aligned16(p->cos) || !aligned16(p->sin))
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "rope: all pointers must be 16-byte aligned");
    auto strides_ok = [](const int64_t* s, const int64_t* dims, int n) {
        for (int i = 0; i < n; ++i)
            if (dims[i] > 1 && (s[i] < 0 || s[i] % 8 != 0)) return false;
        return true;
    };
    const int64_t qd[3] = {p->batch, p->heads_q, p->seqlen};
    const int64_t kd[3] = {p->batch, p->heads_kv, p->seqlen};
    const int64_t cd[2] = {p->batch, p->seqlen};
    if (!strides_ok(p->q_strides, qd, 3) || !strides_ok(p->k_strides, kd, 3) ||
        !strides_ok(p->q_out_strides, qd, 3) || !strides_ok(p->k_out_strides, kd, 3) ||
        !strides_ok(p->cos_sin_strides, cd, 2))
        return fail(FA_SM100_ERR_INVALID_ARGUMENT,
                    "rope: strides must be non-negative multiples of 8 elements (16 bytes)");
    const long long total = (long long)p->batch * p->seqlen * (p->heads_q + p->heads_kv) * (p->headdim / 16);
    if (p->seqlen > (1ll << 30) || p->heads_q + p->heads_kv > (1 << 20) || total > (1ll << 40) ||
        p->batch * p->seqlen > 0x7fffffffll)
        return fail(FA_SM100_ERR_INVALID_ARGUMENT, "problem size out of range");

    DeviceState* st = device_state(p->device);
    if (!st) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "bad device ordinal %d", p->device);
    if (st->error) return fail(st->error, "device init failed: %s", st->msg);
    if (st->cc_major != 10)
        return fail(FA_SM100_ERR_UNSUPPORTED_DEVICE,
                    "fa_sm100 needs an sm_100 (B200) GPU; device %d is sm_%d%d — there is no fallback",
                    p->device, st->cc_major, st->cc_minor);
    int prev = 0;
    cudaGetDevice(&prev);
    if (prev != p->device) cudaSetDevice(p->device);
    struct Restore {
        int prev, cur;
        ~Restore() { if (prev != cur) cudaSetDevice(prev); }
    } restore{prev, p->device};

    fa100::rope::RopeParams kp;
    memset(&kp, 0, sizeof(kp));
    kp.q = p->q; kp.k = p->k; kp.q_out = p->q_out; kp.k_out = p->k_out; kp.cos = p->cos; kp.sin = p->sin;
    kp.batch = int(p->batch); kp.heads_q = int(p->heads_q); kp.heads_kv = int(p->heads_kv);
    kp.seqlen = int(p->seqlen); kp.headdim = int(p->headdim);
    for (int i = 0; i < 3; ++i) {
        kp.q_s[i] = p->q_strides[i]; kp.k_s[i] = p->k_strides[i];
        kp.qo_s[i] = p->q_out_strides[i]; kp.ko_s[i] = p->k_out_strides[i];
    }
    kp.cs_sb = p->batch > 1 ? p->cos_sin_strides[0] : 0;
    kp.cs_ss = p->cos_sin_strides[1];
    kp.total = total;
    // enough CTAs to keep every SM's load queue full (8 per SM), grid-stride over the rest
    long long blocks = (long long)p->batch * p->seqlen;  // one (batch, position) row per loop turn
    const long long cap = 8ll * (st->num_sms > 0 ? st->num_sms : 148);
    if (blocks > cap) blocks = cap;
    if (p->dtype == FA_SM100_BF16)
        fa100::rope::fa_rope_sm100<true><<<dim3((unsigned)blocks), dim3(fa100::rope::kThreads), 0,
                                           static_cast<cudaStream_t>(stream)>>>(kp);
    else
        fa100::rope::fa_rope_sm100<false><<<dim3((unsigned)blocks), dim3(fa100::rope::kThreads), 0,
                                            static_cast<cudaStream_t>(stream)>>>(kp);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(FA_SM100_ERR_CUDA, "rope kernel launch: %s", cudaGetErrorString(e));
    g_launches.fetch_add(1);
    return FA_SM100_OK;
}

extern "C" int fa_sm100_check_device(int device) {
    DeviceState* st = device_state(device);
    if (!st) return fail(FA_SM100_ERR_INVALID_ARGUMENT, "bad device ordinal %d", device);
    if (st->error) return fail(st->error, "device init failed: %s", st->msg);
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(device);
    unsigned int h[4] = {0, 0, 0, 0};
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpy(h, st->status, sizeof(h), cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && h[0] != 0) {
        cudaMemset(st->status, 0, sizeof(h));
        cudaMemset(st->sched, 0, kSchedSlots * 2 * sizeof(int));  // an aborted kernel leaves its slot dirty
    }
    cudaSetDevice(prev);
    if (e != cudaSuccess) return fail(FA_SM100_ERR_CUDA, "device sync: %s", cudaGetErrorString(e));
    if (h[0] != 0)
        return fail(FA_SM100_ERR_KERNEL_ABORT,
                    "in-kernel watchdog fired: wait site %u, block %u, thread %u, blocked-site mask 0x%x",
                    h[0] & 0x7fffffffu, h[1], h[2], h[3]);
    return FA_SM100_OK;
}

// Exposed for fa_sm100_selftest.cu
extern "C" unsigned int* fa_sm100_internal_status(int device) {
    DeviceState* st = device_state(device);
    return (st && !st->error) ? st->status : nullptr;
}
extern "C" int fa_sm100_internal_encode_2d(CUtensorMap* map, int is_bf16, const void* base,
                                           uint64_t rows, uint64_t cols) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {64, 128};
    return encode_map(map, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                      2, base, dims, strides, box, "selftest");
}
extern "C" void fa_sm100_internal_count_launch(void) { g_launches.fetch_add(1); }
// Debug: device buffer of 5 x 4096 int64 that CTA 0 fills with (event << 56 | clock64) records;
// only honoured by FA_SM100_TUNE builds (D=128 bf16 / D=64 fp16).  NULL switches tracing off.
extern "C" void fa_sm100_internal_set_trace(void* device_buffer) {
    g_trace = static_cast<long long*>(device_buffer);
}
extern "C" int fa_sm100_internal_fail(int code, const char* msg) { return fail(code, "%s", msg); }
