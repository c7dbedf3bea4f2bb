// Work tiles of the sm_100a forward kernel: kernel parameters, the decoded tile record, the
// tile order (LPT inside L2-sized sections) for fixed-length batches and for packed
// variable-length batches, and the tiny plan kernel that orders the sequences of a varlen batch.
//
// Replaces the reference's grid mapping (csrc/flash_attention_template.cuh:157-160, 567-583:
// grid = (m-blocks, heads, batch), one CTA per tile, kv-head = q-head / G) with a persistent
// grid that pulls tiles from one global counter.
#pragma once
#include "fa_sm100_ptx.cuh"

namespace fa100 {

constexpr int kBlockM = 128;             // query rows of one M-tile (one TMEM lane each)
constexpr int kBlockN = 128;             // keys per KV block (visited as two 64-key half blocks)
constexpr int kPanelBytes = 128 * 128;   // 128 rows x 64 16-bit elements, 128B-swizzled
constexpr int kSchedDepth = 4;           // decoded tile records in flight inside a CTA
constexpr float kRescaleThreshold = 8.0f;  // lazy rescale: only when a row max grows by > 2^8

struct FwdParams {
    CUtensorMap tm_q;  // 5-D (d, s, g, hkv, b); box (64, rows_per_head, pack_g, 1, 1)
    CUtensorMap tm_k;  // 4-D (d, s, hkv, b);    box (64, 64, 1, 1)
    CUtensorMap tm_v;  // 4-D (d, s, hkv, b);    box (64, 64, 1, 1)
    void* o;
    int64_t o_stride_b, o_stride_h, o_stride_s;  // elements
    int32_t batch, heads_q, heads_kv, group;     // group = heads_q / heads_kv
    int32_t seqlen_q, seqlen_k, headdim;         // fixed-length batches (varlen: per tile record)
    int32_t pack_g;         // q heads packed into one M tile (1 = unpacked, else == group)
    int32_t rows_per_head;  // P
    int32_t num_m_blocks;   // ceil(seqlen_q / (2 P))                       (fixed-length)
    int32_t head_units;     // pack ? heads_kv : heads_q
    int32_t total_heads;    // head_units * batch                           (fixed-length)
    int32_t section_heads;  // heads per L2 section (tile order)            (fixed-length)
    int32_t num_tiles;      // num_m_blocks * total_heads; varlen: read from the plan
    int32_t q_tile_bytes;   // bytes TMA delivers for one Q M-tile (all panels)
    int32_t causal;
    float scale_log2;       // softmax_scale * log2(e)
    WaitCtx wait;           // abort flag + watchdog
    long long* trace;       // debug timeline (kTrace builds only), else nullptr
    int* sched;             // [0] next-tile counter, [1] finished-CTA counter (both zero between launches)
    // packed variable-length batches (nullptr / 0 for fixed-length launches)
    const int32_t* cu_q;    // device, num_seqs + 1 token offsets of q / o
    const int32_t* cu_k;    // device, num_seqs + 1 token offsets of k / v
    const int32_t* plan;    // device, written by varlen_plan_kernel (layout below)
    int32_t num_seqs;
};

// One CTA's view of a work tile (two M-tiles of 128 rows sharing one K/V stream): 16 ints.
struct Tile {
    int mblock, hu, b;   // m-block, head unit (q head or kv head when packed), batch (varlen: 0)
    int kv_head, g_coord;
    int pos0[2];         // first query position of each M-tile inside its sequence
    int nblk[2];         // KV blocks each M-tile visits
    int nmax;
    int seqlen_q, seqlen_k;  // of this tile's sequence
    int q_row0, k_row0;      // first token row of the sequence in q / o and in k / v (varlen)
    int flags;               // kTileVFix
};
constexpr int kTileVFix = 1;  // the last K/V stage holds rows of the NEXT sequence: zero V's tail rows

// Number of 128-wide KV blocks an M-tile starting at query position `tile_pos0` has to visit.
__device__ __forceinline__ int tile_num_blocks(int rows_per_head, int causal, int seqlen_q, int seqlen_k,
                                               int tile_pos0) {
    if (tile_pos0 >= seqlen_q) return 0;
    int kv_end = seqlen_k;
    if (causal) {
        int maxpos = min(tile_pos0 + rows_per_head, seqlen_q) - 1;
        kv_end = min(kv_end, maxpos + (seqlen_k - seqlen_q) + 1);
    }
    return kv_end <= 0 ? 0 : (kv_end + kBlockN - 1) / kBlockN;
}

__device__ __forceinline__ void tile_finish(const FwdParams& p, Tile& t) {
    const bool packed = p.pack_g > 1;
    t.kv_head = packed ? t.hu : t.hu / p.group;
    t.g_coord = packed ? 0 : t.hu % p.group;
    const int P = p.rows_per_head;
    t.pos0[0] = t.mblock * 2 * P;
    t.pos0[1] = t.pos0[0] + P;
    t.nblk[0] = tile_num_blocks(P, p.causal, t.seqlen_q, t.seqlen_k, t.pos0[0]);
    t.nblk[1] = tile_num_blocks(P, p.causal, t.seqlen_q, t.seqlen_k, t.pos0[1]);
    t.nmax = max(t.nblk[0], t.nblk[1]);
}

// Fixed-length order.  Linear index L -> (section, m-rank, head): sections of `section_heads`
// consecutive (batch, head) units whose K/V fit in L2 together; inside a section the m-blocks go
// heaviest (last) first and the heads vary fastest, so CTAs running at the same time read the
// same few K/V heads.
__device__ __forceinline__ void decode_tile(const FwdParams& p, int L, Tile& t) {
    const int nm = p.num_m_blocks;
    const int sec_tiles = p.section_heads * nm;
    const int sec = L / sec_tiles;
    const int rem = L - sec * sec_tiles;
    const int h0 = sec * p.section_heads;
    const int nh = min(p.section_heads, p.total_heads - h0);
    const int mrank = rem / nh;
    const int hflat = h0 + (rem - mrank * nh);
    t.mblock = nm - 1 - mrank;
    t.hu = hflat % p.head_units;
    t.b = hflat / p.head_units;
    t.seqlen_q = p.seqlen_q;
    t.seqlen_k = p.seqlen_k;
    t.q_row0 = 0;
    t.k_row0 = 0;
    t.flags = 0;
    tile_finish(p, t);
}

// ---- packed variable-length batches --------------------------------------------------------------
// plan (int32): [0] total tiles, [1] n, [2 .. 2+n) sequence index by descending key count (LPT:
// the longest sequences' tiles are handed out first), [2+n .. 3+2n) first tile of each sorted
// position (exclusive prefix sum of m-blocks x head_units) with the total at the end.
// Inside a sequence the order is the fixed-length one: heaviest m-block first, heads fastest.
__host__ __device__ constexpr int64_t varlen_plan_ints(int64_t n) { return 3 + 2 * n; }

// Tile indices reach a CTA in increasing order, so the sorted position only ever moves forward.
struct VarlenCursor {
    int pos = 0, start = 0, end = 0;
    bool primed = false;
};

__device__ __forceinline__ void decode_tile_varlen(const FwdParams& p, int L, VarlenCursor& c, Tile& t) {
    const int n = p.num_seqs;
    const int32_t* sorted = p.plan + 2;
    const int32_t* first = p.plan + 2 + n;
    if (!c.primed) {
        c.primed = true;
        c.pos = 0;
        c.start = 0;
        c.end = first[1];
    }
    while (L >= c.end) {  // L < total tiles: terminates before pos reaches n
        ++c.pos;
        c.start = c.end;
        c.end = first[c.pos + 1];
    }
    const int seq = sorted[c.pos];
    const int rem = L - c.start;
    const int mrank = rem / p.head_units;
    t.hu = rem - mrank * p.head_units;
    t.b = 0;
    t.q_row0 = p.cu_q[seq];
    t.k_row0 = p.cu_k[seq];
    t.seqlen_q = p.cu_q[seq + 1] - t.q_row0;
    t.seqlen_k = p.cu_k[seq + 1] - t.k_row0;
    const int nm = (t.seqlen_q + 2 * p.rows_per_head - 1) / (2 * p.rows_per_head);
    t.mblock = nm - 1 - mrank;
    tile_finish(p, t);
    // Rows behind the sequence's last key belong to the next sequence (TMA only zero-fills past
    // the end of the packed tensor).  K: their scores are masked.  V: P is exactly 0 there, but
    // 0 x inf/NaN would still poison the row, so those V rows are zeroed in shared memory.
    t.flags = (t.nmax * kBlockN > t.seqlen_k) ? kTileVFix : 0;
}

constexpr int kPlanThreads = 1024;
// One CTA.  Sequences are ranked by key count (ties by index); batches of more than
// kPlanSortMax sequences keep their given order (still correct, only the tail balance suffers).
constexpr int kPlanSortMax = 2048;
__global__ void __launch_bounds__(kPlanThreads, 1)
varlen_plan_kernel(const int32_t* __restrict__ cu_q, const int32_t* __restrict__ cu_k, int n,
                   int rows_per_tile, int head_units, int32_t* __restrict__ plan) {
    int32_t* sorted = plan + 2;
    int32_t* first = plan + 2 + n;
    const int tid = threadIdx.x;
    __shared__ int warp_sums[kPlanThreads / 32];
    __shared__ int carry;
    __shared__ int keys[kPlanSortMax];
    const bool sort = n <= kPlanSortMax;
    if (sort) {
        for (int i = tid; i < n; i += kPlanThreads) keys[i] = cu_k[i + 1] - cu_k[i];
        __syncthreads();
    }
    for (int i = tid; i < n; i += kPlanThreads) {
        int rank = i;
        if (sort) {
            const int ki = keys[i];
            rank = 0;
            for (int j = 0; j < n; ++j) {
                const int kj = keys[j];
                rank += (kj > ki || (kj == ki && j < i)) ? 1 : 0;
            }
        }
        sorted[rank] = i;
    }
    __syncthreads();
    if (tid == 0) carry = 0;
    __syncthreads();
    // exclusive prefix sum of tiles per sorted position, kPlanThreads entries per turn
    for (int base = 0; base < n; base += kPlanThreads) {
        const int i = base + tid;
        int tiles = 0;
        if (i < n) {
            const int s = sorted[i];
            const int sq = cu_q[s + 1] - cu_q[s];
            tiles = ((sq + rows_per_tile - 1) / rows_per_tile) * head_units;
        }
        int incl = tiles;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int up = __shfl_up_sync(0xffffffffu, incl, d);
            if ((tid & 31) >= d) incl += up;
        }
        if ((tid & 31) == 31) warp_sums[tid >> 5] = incl;
        __syncthreads();
        if (tid < 32) {
            int ws = warp_sums[tid];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int up = __shfl_up_sync(0xffffffffu, ws, d);
                if (tid >= d) ws += up;
            }
            warp_sums[tid] = ws;
        }
        __syncthreads();
        const int before = carry + ((tid >> 5) ? warp_sums[(tid >> 5) - 1] : 0) + incl - tiles;
        if (i < n) first[i] = before;
        __syncthreads();
        if (tid == kPlanThreads - 1) carry = before + tiles;
        __syncthreads();
    }
    if (tid == 0) {
        first[n] = carry;
        plan[0] = carry;
        plan[1] = n;
    }
}

// Debug timeline: CTA 0 records (event code << 56 | clock64) per role; see tools/trace_view.py.
constexpr int kTraceRoles = 5;
constexpr int kTraceMax = 4096;
template <bool kTrace>
struct Tracer {
    long long* buf;
    int n;
    __device__ __forceinline__ void init(const FwdParams& p, int role, bool on) {
        if constexpr (kTrace) {
            buf = (on && p.trace && blockIdx.x == 0) ? p.trace + role * kTraceMax : nullptr;
            n = 0;
        }
    }
    __device__ __forceinline__ void ev(int code) {
        if constexpr (kTrace) {
            if (buf && n < kTraceMax) buf[n++] = (static_cast<long long>(code) << 56) | (clock64_now() & 0xffffffffffffffll);
        }
    }
};

}  // namespace fa100
