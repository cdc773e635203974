// K2 (multistart / POMO) on the 5th-generation tensor cores.
//
// With S starts per instance the decode step IS three GEMMs that share the instance's tables
// (rl4co/models/zoo/am/decoder.py:176-200, rl4co/models/nn/attention.py:263-309):
//     scores_j [S x N]  = q_j [S x 16] K_j^T [16 x N]        per head j (8 of them)
//     glimpse_j [S x 16] = P_j [S x N] V_j [N x 16]          per head j
//     logits [S x N]    = o [S x 128] L'^T [128 x N]         (project_out folded into L', as in the SIMT kernels)
// One CTA per SM works on one instance for its whole trajectory:
//   * the S <= 128 starts are the M rows = the 128 TMEM LANES; consumer thread = start = lane, so the masked softmax,
//     the pointer softmax, arg-max / sampling and the environment transition of a rollout need NO cross-thread
//     communication at all;
//   * the three tables stay resident in shared memory as the B operands (K-major, 128-byte swizzled rows);
//   * every A operand - the query, the softmax weights P, the normalised glimpse - is written by its owner thread
//     straight into TENSOR MEMORY (tcgen05.st) and read from there by tcgen05.mma; accumulators (scores, glimpse,
//     logits) come back with tcgen05.ld.
// Arithmetic: fp32 accuracy from an FP16 two-term split, x = hi + lo with hi = fp16(x), lo = fp16(x - hi)
// (22 significant bits), three kind::f16 MMAs per product (lo*hi + hi*lo + hi*hi, fp32 accumulation).  The TF32 split
// of the projection kernel would need 344 KB of shared memory for the tables; FP16 halves that (176 KB) and runs at
// twice the TF32 rate.  Values below 2^-14 lose low-order bits of `lo` (absolute error <= 3e-8 per element).
//
// Pipeline of one step (MMA warp | 2 consumer groups of 4 warps; group g owns heads 4g..4g+3 and score buffer g):
//   consumers  q = scale (gctx + ctx rows + capacity column)  -> TMEM [QH|QL]                      -> q_ready
//   MMA        scores(0), scores(1)                           -> S0, S1                           -> s_full[b]
//   consumers  head j: max, P = 2^(s - max) (masked: 0), l_j  -> P_hi over S, P_lo beside it       -> p_full[b]
//   MMA        O[b] = P_j V_j ; scores(j + 2) -> S[b]                                             -> o_full[b], s_full[b]
//   consumers  glimpse_j = O[b] / l_j                         -> TMEM over the dead q_j columns
//   ... after head 7                                                                              -> o_ready
//   MMA        logits = o L'^T                                -> S0                               -> lg_full
//   consumers  clip, mask, log-softmax ingredients, arg-max / inverse-CDF sample / given action, env transition
//
// Environments: tsp and cvrp (the multistart environments of BASELINE.json), standard logit post-processing
// (tanh clipping > 0, temperature 1, masked logits); everything else stays on am_rollout_rf.cu.
// Reference semantics reproduced: utils/decoding.py:265-313 (forced start node, log-prob 0), :141-191, :371-397,
// envs/routing/tsp/env.py:67-93,157-163, envs/routing/cvrp/env.py:73-160, utils/ops.py:139-163.
#include <cuda_fp16.h>

#include "rollout_util.cuh"
#include "tc_util.cuh"

namespace rl4co {
namespace mstc {
using namespace tcu;

constexpr int E = RL4CO_EMBED_DIM, H = RL4CO_NUM_HEADS, D = 16;
constexpr int NK = 112;                                 // node columns of the score / logit tiles (multiple of 16)
constexpr int kGroupWarps = 4, kGroups = 2;             // a consumer group covers the 128 TMEM lanes once
constexpr int kConsumerWarps = kGroups * kGroupWarps;
constexpr int kThreads = (1 + kConsumerWarps) * 32;     // warp 0: MMA issue; warps 1-4 / 5-8: thread = (start, group)
constexpr int HG = H / kGroups;                         // heads per group
constexpr int NH = NK / kGroups;                        // logit columns per group (56)
constexpr float kScoreScale = 0.25f * 1.4426950408889634f;   // 1/sqrt(16) * log2(e)
constexpr float kInvSqrtE = 0.08838834764831845f;
constexpr float kLog2e = 1.4426950408889634f;

// ---- shared memory: B-operand images, rows of 128 B = 64 fp16, 8-row groups of 1024 B, 128B swizzle
constexpr int KT_TILE = NK * 128;                       // K / L: [112 nodes][64 dims]   (2 tiles per image)
constexpr int VT_TILE = 128 * 128;                      // V^T:   [128 dims][64 nodes]   (2 tiles per image)
constexpr int OFF_KH = 0, OFF_KL = 2 * KT_TILE, OFF_LH = 4 * KT_TILE, OFF_LL = 6 * KT_TILE;
constexpr int OFF_VH = 8 * KT_TILE, OFF_VL = OFF_VH + 2 * VT_TILE, OFF_MISC = OFF_VL + 2 * VT_TILE;   // 180 224
struct Misc {
    float dem[128];         // cvrp: demand of node n (0 for the depot and the pad nodes)
    float2 locs[128];
    float wc[128];          // cvrp: remaining-capacity column of project_context, pre-scaled
    float gq[128];          // graph context, pre-scaled (0 without one)
    float stage[kConsumerWarps][32 * 36];   // query build: 32 starts x 32 dims per warp and pass, rows padded to 36 floats
    float4 xch[kGroups][128];   // logit phase: each group's partial (sum of e, best logit, its node, its e) per start
    uint64_t bars[12];
    uint32_t tmem_slot;
    int stop;               // written by one consumer before q_ready: every rollout of the instance is done
};
constexpr int SMEM_BYTES = 1024 + OFF_MISC + (int)sizeof(Misc);
enum { B_QREADY = 0, B_SFULL = 1, B_PFULL = 3, B_OFULL = 5, B_OREADY = 7, B_LGFULL = 8 };

// ---- tensor memory (512 columns x 128 lanes)
constexpr uint32_t C_QH = 0, C_QL = 64;                 // query, later the normalised glimpse: fp16 pairs, hi | lo
constexpr uint32_t C_S0 = 128, C_PL0 = 240, C_S1 = 296, C_PL1 = 408;   // scores -> P_hi in place (56 cols), P_lo (56 cols)
constexpr uint32_t C_O0 = 464, C_O1 = 480;              // P V accumulators, 16 columns each
constexpr uint32_t C_LG = C_S0;                         // logits reuse the first score buffer

// instruction descriptor (cute/arch/mma_sm100_desc.hpp): D = F32, A = B = F16 (format 0), K-major, N/8, M/16
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n) {
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
constexpr uint32_t kIdescS = make_idesc_f16(128, NK);
constexpr uint32_t kIdescO = make_idesc_f16(128, D);

// D[tmem] (+)= A[tmem, fp16 pairs per column] * B[smem descriptor]
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_lo32, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 db;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "r"(tmem_a), "r"(b_lo32), "r"(idesc), "r"(accumulate), "r"(kDescHi)
        : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(v[0]),
                 "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float ex2f(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float rcpf(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// x0, x1 -> fp16 pair hi (x0 in the low half = the even K element) and the pair of residuals lo
__device__ __forceinline__ void split2(float x0, float x1, uint32_t &hi, uint32_t &lo) {
    const __half2 h = __floats2half2_rn(x0, x1);
    const float2 back = __half22float2(h);
    const __half2 l = __floats2half2_rn(x0 - back.x, x1 - back.y);
    hi = *reinterpret_cast<const uint32_t *>(&h);
    lo = *reinterpret_cast<const uint32_t *>(&l);
}
// barrier among the 256 consumer threads with an AND reduction of a predicate
__device__ __forceinline__ bool consumers_all(bool pred) {
    uint32_t r;
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "setp.ne.b32 q, %1, 0;\n\t"
        "bar.red.and.pred p, 1, 256, q;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n"
        : "=r"(r)
        : "r"((uint32_t)pred)
        : "memory");
    return r != 0;
}

// ---- per-instance staging of the tables: fp32 rows in global memory -> fp16 hi / lo operand images ---------------
// K and L': element (node n, dim d) -> tile d >> 6, row n, 16-byte chunk ((d & 63) >> 3) ^ (n & 7), half d & 7
__device__ __forceinline__ void stage_rows(unsigned char *img_hi, unsigned char *img_lo, const float *rows, int N, int tid) {
    for (int item = tid; item < N * 16; item += kThreads) {
        const int n = item >> 4, ch = item & 15;            // 8 dims: one 16-byte chunk of each image
        const float4 a = __ldg(reinterpret_cast<const float4 *>(rows + (size_t)n * E + ch * 8));
        const float4 b = __ldg(reinterpret_cast<const float4 *>(rows + (size_t)n * E + ch * 8) + 1);
        uint4 hi, lo;
        split2(a.x, a.y, hi.x, lo.x);
        split2(a.z, a.w, hi.y, lo.y);
        split2(b.x, b.y, hi.z, lo.z);
        split2(b.z, b.w, hi.w, lo.w);
        const uint32_t off = (uint32_t)((ch >> 3) * KT_TILE + n * 128 + (((ch & 7) ^ (n & 7)) << 4));
        *reinterpret_cast<uint4 *>(img_hi + off) = hi;
        *reinterpret_cast<uint4 *>(img_lo + off) = lo;
    }
}
// V^T: element (dim d, node n) -> tile n >> 6, row d, chunk ((n & 63) >> 3) ^ (d & 7), half n & 7
__device__ __forceinline__ void stage_transposed(unsigned char *img_hi, unsigned char *img_lo, const float *rows, int N, int tid) {
    for (int item = tid; item < N * 32; item += kThreads) {
        const int n = item >> 5, d0 = (item & 31) * 4;      // 4 dims of node n
        const float4 v = __ldg(reinterpret_cast<const float4 *>(rows + (size_t)n * E + d0));
        const float x[4] = {v.x, v.y, v.z, v.w};
        const uint32_t col = (uint32_t)((n >> 6) * VT_TILE + (n & 7) * 2);
        const uint32_t chunk = (uint32_t)((n & 63) >> 3);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int d = d0 + i;
            const __half h = __float2half_rn(x[i]);
            const __half l = __float2half_rn(x[i] - __half2float(h));
            const uint32_t off = col + (uint32_t)d * 128u + ((chunk ^ (uint32_t)(d & 7)) << 4);
            *reinterpret_cast<__half *>(img_hi + off) = h;
            *reinterpret_cast<__half *>(img_lo + off) = l;
        }
    }
}

__device__ __forceinline__ void tmem_ld32_nw(uint32_t taddr, uint32_t *v) {       // no wait: pair with tmem_ld_wait()
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16_nw(uint32_t taddr, uint32_t *v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld8_nw(uint32_t taddr, uint32_t *v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// the NH = 56 logit columns of one consumer group
__device__ __forceinline__ void tmem_ld_half(uint32_t taddr, uint32_t (&v)[NH]) {
    tmem_ld32_nw(taddr, v);
    tmem_ld16_nw(taddr + 32, v + 32);
    tmem_ld8_nw(taddr + 48, v + 48);
    tmem_ld_wait();
}
__device__ __forceinline__ void consumers_sync() { asm volatile("bar.sync 2, 256;" ::: "memory"); }
// all NK columns of a score / logit tile of this thread's lane
__device__ __forceinline__ void tmem_ld_row(uint32_t taddr, uint32_t (&v)[NK]) {
    tmem_ld32_nw(taddr, v);
    tmem_ld32_nw(taddr + 32, v + 32);
    tmem_ld32_nw(taddr + 64, v + 64);
    tmem_ld16_nw(taddr + 96, v + 96);
    tmem_ld_wait();
}

// -DMS_TRACE (tools/trace_ms_tc.py): the consumer thread of start 0 in CTA 0 stamps clock64() during its first instance
// into a.logp_all, reused as an int64 [T][16] buffer: 0 step start, 1 query stored, 2 first scores arrived, 3 heads done,
// 4 last glimpse stored, 5 logits arrived, 6 selection done, 7 step end; 8 / 9 / 10: cycles spent waiting for scores /
// for P V accumulators / inside the softmax arithmetic, summed over the 8 heads
#ifdef MS_TRACE
#define MS_STAMP(k) do { if (trace_on) trace[(size_t)t * 16 + (k)] = clock64(); } while (0)
#define MS_ACC(k, expr) do { const long long c0__ = clock64(); expr; if (trace_on) trace[(size_t)t * 16 + (k)] += clock64() - c0__; } while (0)
#else
#define MS_STAMP(k) do { } while (0)
#define MS_ACC(k, expr) do { expr; } while (0)
#endif

typedef unsigned long long u64;
// node sets as four 32-bit words; with a compile-time n this is one LOP3 + ISETP
struct Bits { unsigned w[4]; };
__device__ __forceinline__ bool bit_at(const Bits &m, int n) { return (m.w[n >> 5] >> (n & 31)) & 1u; }
__device__ __forceinline__ Bits make_bits(u64 lo, u64 hi) {
    Bits m;
    m.w[0] = (unsigned)lo; m.w[1] = (unsigned)(lo >> 32); m.w[2] = (unsigned)hi; m.w[3] = (unsigned)(hi >> 32);
    return m;
}

// registers are allocated to warps in groups of four: 9 warps are budgeted as 12, i.e. 168 registers per thread - the
// softmax and the sampling scan therefore work on 64 + 48 columns at a time instead of holding a whole row
template <int ENV, int MODE>
__global__ void __launch_bounds__(kThreads, 1)
am_rollout_ms_tc_kernel(const rl4co_am_rollout_args a) {
    constexpr bool CVRP = ENV == RL4CO_ENV_CVRP;
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    Misc &ms = *reinterpret_cast<Misc *>(smem + OFF_MISC);
    uint64_t *bars = ms.bars;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int N = a.N, T = a.T, S = a.S;
    const int64_t Mrows = a.B * (int64_t)N;
    const float clip = a.tanh_clipping;

    // pad rows (nodes >= N) of every operand image must read as zero for the whole launch
    for (int off = tid * 16; off < OFF_MISC; off += kThreads * 16) *reinterpret_cast<uint4 *>(smem + off) = make_uint4(0u, 0u, 0u, 0u);
    if (tid == 0) {
        mbar_init(&bars[B_QREADY], kConsumerWarps);         // every consumer warp stores its part of the query
        mbar_init(&bars[B_SFULL], 1);      mbar_init(&bars[B_SFULL + 1], 1);
        mbar_init(&bars[B_PFULL], kGroupWarps);     mbar_init(&bars[B_PFULL + 1], kGroupWarps);   // per group
        mbar_init(&bars[B_OFULL], 1);      mbar_init(&bars[B_OFULL + 1], 1);
        mbar_init(&bars[B_OREADY], kConsumerWarps);
        mbar_init(&bars[B_LGFULL], 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ms.tmem_slot)), "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid < 128) ms.wc[tid] = (CVRP && a.wcap) ? a.wcap[tid] * kScoreScale : 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ms.tmem_slot;

    // running phase parities of the mbarriers (bit b = buffer b)
    unsigned ph_q = 0, ph_p = 0, ph_or = 0, ph_lg = 0;

    for (int64_t b = blockIdx.x; b < a.B; b += gridDim.x) {
        __syncthreads();                                        // the previous instance is done with shared memory
        // ---------------------------------------------------------------- tables -> operand images
        stage_rows(smem + OFF_KH, smem + OFF_KL, a.tables + ((size_t)0 * Mrows + b * (int64_t)N) * E, N, tid);
        stage_rows(smem + OFF_LH, smem + OFF_LL, a.tables + ((size_t)2 * Mrows + b * (int64_t)N) * E, N, tid);
        stage_transposed(smem + OFF_VH, smem + OFF_VL, a.tables + ((size_t)1 * Mrows + b * (int64_t)N) * E, N, tid);
        if (tid < 128) {
            ms.dem[tid] = (CVRP && tid >= 1 && tid < N) ? a.demand[b * (int64_t)(N - 1) + tid - 1] : 0.f;
            ms.gq[tid] = a.gctx ? a.gctx[b * E + tid] * kScoreScale : 0.f;
            ms.locs[tid] = tid < N ? reinterpret_cast<const float2 *>(a.locs + b * (int64_t)N * 2)[tid] : make_float2(0.f, 0.f);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic writes -> async proxy (UMMA reads)
        __syncthreads();

        if (warp == 0) {
            // ============================================================ MMA issue
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
            const uint32_t d0 = (uint32_t)smem_desc(smem_u32(smem));
            const uint32_t kh = d0 + (OFF_KH >> 4), kl = d0 + (OFF_KL >> 4), lh = d0 + (OFF_LH >> 4), ll = d0 + (OFF_LL >> 4);
            const uint32_t vh = d0 + (OFF_VH >> 4), vl = d0 + (OFF_VL >> 4);
            auto scores = [&](int j) {                          // head j of group j / HG: S[group] = q_j K_j^T, three split terms
                const int g = j / HG;
                const uint32_t boff = (uint32_t)(((j >> 2) * KT_TILE + (j & 3) * 32) >> 4);
                const uint32_t dcol = tmem_u + (g ? C_S1 : C_S0);
                const uint32_t qh = tmem_u + C_QH + 8u * j, ql = tmem_u + C_QL + 8u * j;
                umma_f16_ts(dcol, ql, kh + boff, kIdescS, 0u);
                umma_f16_ts(dcol, qh, kl + boff, kIdescS, 1u);
                umma_f16_ts(dcol, qh, kh + boff, kIdescS, 1u);
                umma_commit(&bars[B_SFULL + g]);
            };
            for (;;) {
                mbar_wait_bounded(&bars[B_QREADY], ph_q);
                ph_q ^= 1u;
                tc_fence_after();
                if (*reinterpret_cast<volatile int *>(&ms.stop)) break;
                if (elect_one()) {
                    scores(0);
                    scores(HG);
                }
                __syncwarp();
#pragma unroll 1
                for (int it = 0; it < H; ++it) {                // the two groups alternate: head jj of group 0, head jj of group 1
                    const int g = it & 1, j = g * HG + (it >> 1);
                    mbar_wait_bounded(&bars[B_PFULL + g], (ph_p >> g) & 1u);
                    ph_p ^= 1u << g;
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t o = tmem_u + (g ? C_O1 : C_O0);
                        const uint32_t p_hi = tmem_u + (g ? C_S1 : C_S0), p_lo = tmem_u + (g ? C_PL1 : C_PL0);
#pragma unroll
                        for (int ks = 0; ks < NK / 16; ++ks) {  // 7 k-steps of 16 nodes
                            const uint32_t voff = (uint32_t)(((ks >> 2) * VT_TILE + j * 16 * 128 + (ks & 3) * 32) >> 4);
                            umma_f16_ts(o, p_lo + 8u * ks, vh + voff, kIdescO, (unsigned)(ks != 0));
                            umma_f16_ts(o, p_hi + 8u * ks, vl + voff, kIdescO, 1u);
                            umma_f16_ts(o, p_hi + 8u * ks, vh + voff, kIdescO, 1u);
                        }
                        umma_commit(&bars[B_OFULL + g]);
                        if ((it >> 1) + 1 < HG) scores(j + 1);  // executes after the P V above: the group's S is free again
                    }
                    __syncwarp();
                }
                mbar_wait_bounded(&bars[B_OREADY], ph_or);
                ph_or ^= 1u;
                tc_fence_after();
                if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < E / 16; ++ks) {       // 8 k-steps of 16 dims
                        const uint32_t loff = (uint32_t)(((ks >> 2) * KT_TILE + (ks & 3) * 32) >> 4);
                        const uint32_t oh = tmem_u + C_QH + 8u * ks, ol = tmem_u + C_QL + 8u * ks;
                        umma_f16_ts(tmem_u + C_LG, ol, lh + loff, kIdescS, (unsigned)(ks != 0));
                        umma_f16_ts(tmem_u + C_LG, oh, ll + loff, kIdescS, 1u);
                        umma_f16_ts(tmem_u + C_LG, oh, lh + loff, kIdescS, 1u);
                    }
                    umma_commit(&bars[B_LGFULL]);
                }
                __syncwarp();
            }
        } else {
            // ============================================================ consumers: thread = (start = TMEM lane, group)
            // Both groups cover all 128 lanes; group g owns heads HG g .. HG g + 3 (query dims 64 g .. 64 g + 63, score
            // buffer g) and the logit columns NH g .. NH g + 55.  The environment state of a start is replicated in its
            // two threads (both apply every transition); group 0 writes the outputs.
            const int grp = (warp - 1) / kGroupWarps;
            const int quarter = warp & 3;                       // the TMEM lanes a warp may touch: 32 (warp % 4) ...
            const int sidx = 32 * quarter + lane;
            const uint32_t tl = tmem_base + ((uint32_t)(32 * quarter) << 16);
            const bool active = sidx < S;
            const bool writer = active && grp == 0;
            const int64_t r = (int64_t)sidx * a.B + b;          // rollout index, start-major (utils/ops.py:11-34)
            const u64 own_lo = N >= 64 ? ~0ull : ((1ull << N) - 1ull);
            const u64 own_hi = N <= 64 ? 0ull : ((1ull << (N - 64)) - 1ull);
            u64 vis_lo = 0ull, vis_hi = 0ull, al_lo = own_lo, al_hi = own_hi;
            const float cap = CVRP ? a.vehicle_capacity[b] : 0.f;
            float used = 0.f, len = 0.f, llsum = 0.f;
            float2 prev = ms.locs[0];
            int cur = 0, first = 0, t = 0, status = 0;
            bool done = !active;
            unsigned ph_sg = 0, ph_og = 0;                      // parities of this group's s_full / o_full
            const uint32_t sb = tl + (grp ? C_S1 : C_S0), plb = tl + (grp ? C_PL1 : C_PL0), ob = tl + (grp ? C_O1 : C_O0);
#ifdef MS_TRACE
            long long *trace = reinterpret_cast<long long *>(a.logp_all);
            const bool trace_on = trace != nullptr && blockIdx.x == 0 && sidx == 0 && grp == 0 && b == (int64_t)blockIdx.x;
#endif

            auto cvrp_mask = [&]() {                            // cvrp/env.py:137-149
                u64 f_lo = 0ull, f_hi = 0ull;
#pragma unroll 1
                for (int g = 0; g < 8; ++g) {                   // 16 nodes per round, demands broadcast from shared memory
                    unsigned bits = 0u;
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4) {
                        const float4 d = *reinterpret_cast<const float4 *>(&ms.dem[16 * g + 4 * i4]);
                        bits |= (__fadd_rn(d.x, used) > cap ? 0u : 1u) << (4 * i4);
                        bits |= (__fadd_rn(d.y, used) > cap ? 0u : 2u) << (4 * i4);
                        bits |= (__fadd_rn(d.z, used) > cap ? 0u : 4u) << (4 * i4);
                        bits |= (__fadd_rn(d.w, used) > cap ? 0u : 8u) << (4 * i4);
                    }
                    if (g < 4) f_lo |= (u64)bits << (16 * g);
                    else f_hi |= (u64)bits << (16 * (g - 4));
                }
                f_lo &= own_lo & ~vis_lo & ~1ull;
                f_hi &= own_hi & ~vis_hi;
                if (!(cur == 0 && (f_lo | f_hi) != 0ull)) f_lo |= 1ull;     // the depot, unless we are on it with customers left
                al_lo = f_lo;
                al_hi = f_hi;
            };
            auto apply = [&](int act) {                         // env transition of this rollout
                const float2 p = ms.locs[act];
                if (act < 64) vis_lo |= 1ull << act; else vis_hi |= 1ull << (act - 64);
                if (CVRP) {                                     // cvrp/env.py:73-105
                    len += leg(prev.x, prev.y, p.x, p.y);
                    const float sel = ms.dem[act == 0 ? 1 : act];
                    used = __fmul_rn(__fadd_rn(used, sel), act != 0 ? 1.f : 0.f);
                    cur = act;
                    cvrp_mask();
                    done = vis_lo == own_lo && vis_hi == own_hi;
                } else {                                        // tsp/env.py:67-93
                    if (t == 0) first = act; else len += leg(prev.x, prev.y, p.x, p.y);
                    cur = act;
                    al_lo = own_lo & ~vis_lo;
                    al_hi = own_hi & ~vis_hi;
                    done = t + 1 >= N;
                }
                prev = p;
            };
            if (active) {                                       // forced POMO start node, log-prob 0 (decoding.py:287-305)
                const int a0 = sidx % a.num_loc + (CVRP ? 1 : 0);
                if (writer) {
                    a.actions[(size_t)r * T] = a0;
                    if (a.logp) a.logp[(size_t)r * T] = 0.f;
                }
                apply(a0);
                t = 1;
            }

            for (;;) {
                const bool fin = done || t >= T;
                const bool stop = consumers_all(fin);
                if (warp == 1 && lane == 0) ms.stop = stop ? 1 : 0;
                // masks of this step; finished / unused lanes keep one live column so that their rows stay finite
                const u64 mk_lo = fin ? 1ull : al_lo, mk_hi = fin ? 0ull : al_hi;
                const Bits mk = make_bits(mk_lo, mk_hi);
                MS_STAMP(0);
                if (!stop) {
                    // ---- this group's half of the query: scale * (gctx + ctx(cur) [+ ctx_first(first)]) [+ scaled capacity
                    //      column * remaining capacity]; every load is issued before the first use
                    // The rows are gathered by the WARP, not by the thread: 8 lanes fetch one start's 32 dims as a full
                    // 128-byte line (4 starts per instruction) into a padded staging tile, from which every thread takes
                    // its own row - 64 line requests per warp and step instead of 512 thread-private ones.
                    const float *ctx_tab = a.tables + ((size_t)3 * Mrows + b * (int64_t)N) * E + 64 * grp + 4 * (lane & 7);
                    const float *first_tab = a.tables + ((size_t)4 * Mrows + b * (int64_t)N) * E + 64 * grp + 4 * (lane & 7);
                    const float rem = CVRP ? __fsub_rn(cap, used) : 0.f;
                    const float qs = fin ? 0.f : 1.f;           // finished / unused lanes: a zero query
                    float *stg = ms.stage[warp - 1];
                    float4 g[2][8];                             // every load of both passes is in flight before the first use
#pragma unroll
                    for (int it = 0; it < 8; ++it) {
                        const int src = 4 * it + (lane >> 3);
                        const int cur_s = __shfl_sync(0xffffffffu, cur, src);
                        const int fin_s = __shfl_sync(0xffffffffu, (int)fin, src);
                        const int first_s = CVRP ? 0 : __shfl_sync(0xffffffffu, first, src);
#pragma unroll
                        for (int pass = 0; pass < 2; ++pass) {
                            g[pass][it] = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (!fin_s) g[pass][it] = __ldg(reinterpret_cast<const float4 *>(ctx_tab + (size_t)cur_s * E + 32 * pass));
                        }
                        if (!CVRP && !fin_s) {
#pragma unroll
                            for (int pass = 0; pass < 2; ++pass) {
                                const float4 f = __ldg(reinterpret_cast<const float4 *>(first_tab + (size_t)first_s * E + 32 * pass));
                                g[pass][it].x += f.x; g[pass][it].y += f.y; g[pass][it].z += f.z; g[pass][it].w += f.w;
                            }
                        }
                    }
                    MS_STAMP(11);
#pragma unroll
                    for (int pass = 0; pass < 2; ++pass) {      // 32 dims per pass through the staging tile
#pragma unroll
                        for (int it = 0; it < 8; ++it)
                            *reinterpret_cast<float4 *>(stg + (4 * it + (lane >> 3)) * 36 + 4 * (lane & 7)) = g[pass][it];
                        __syncwarp();
                        if (pass == 0) MS_STAMP(12);
                        uint32_t hi[16], lo[16];
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const int dl = 32 * pass + 4 * k;   // dim within this group's 64
                            float4 v = *reinterpret_cast<const float4 *>(stg + lane * 36 + 4 * k);
                            const float4 gq = *reinterpret_cast<const float4 *>(&ms.gq[64 * grp + dl]);
                            v.x = fmaf(v.x, kScoreScale, gq.x); v.y = fmaf(v.y, kScoreScale, gq.y);
                            v.z = fmaf(v.z, kScoreScale, gq.z); v.w = fmaf(v.w, kScoreScale, gq.w);
                            if (CVRP) {
                                const float4 w = *reinterpret_cast<const float4 *>(&ms.wc[64 * grp + dl]);
                                v.x = fmaf(w.x, rem, v.x); v.y = fmaf(w.y, rem, v.y);
                                v.z = fmaf(w.z, rem, v.z); v.w = fmaf(w.w, rem, v.w);
                            }
                            split2(v.x * qs, v.y * qs, hi[2 * k], lo[2 * k]);
                            split2(v.z * qs, v.w * qs, hi[2 * k + 1], lo[2 * k + 1]);
                        }
                        __syncwarp();                           // the tile is rewritten by the next pass
                        tmem_st16(tl + C_QH + 32u * grp + 16u * pass, hi);
                        tmem_st16(tl + C_QL + 32u * grp + 16u * pass, lo);
                    }
                    MS_STAMP(13);
                }
                tmem_st_wait();
                MS_STAMP(14);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[B_QREADY]);
                if (stop) break;
                MS_STAMP(1);

                // ---- glimpse: masked softmax of this group's heads on the thread's row; the normalised glimpse goes back
                //      to tensor memory, over the dead query columns of the same head, as the A operand of the logits
#pragma unroll 1
                for (int jj = 0; jj < HG; ++jj) {
                    const int j = HG * grp + jj;
                    MS_ACC(8, mbar_wait_bounded(&bars[B_SFULL + grp], ph_sg));
                    ph_sg ^= 1u;
                    tc_fence_after();
                    if (jj == 0) MS_STAMP(2);
#ifdef MS_TRACE
                    const long long soft0 = clock64();
#endif
                    // ONE read of the row: tcgen05.ld moves 64 B/clk per SM (a 128 x 112 fp32 tile = 900 clk), so the masked
                    // scores stay in registers between the max pass and the exp pass (cold per-rollout state spills instead)
                    uint32_t sv[NK];
                    tmem_ld_row(sb, sv);
                    float mx = -INFINITY;
#pragma unroll
                    for (int n = 0; n < NK; ++n) {              // masked scores, row maximum (attention.py:289-306)
                        const float s = bit_at(mk, n) ? __uint_as_float(sv[n]) : -INFINITY;
                        sv[n] = __float_as_uint(s);
                        mx = fmaxf(mx, s);
                    }
                    float l = 0.f;
#pragma unroll
                    for (int c = 0; c < NK / 16; ++c) {         // P = 2^(s - max): hi over the scores, lo beside them
                        uint32_t hi[8], lo[8];
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const float p0 = ex2f(__uint_as_float(sv[16 * c + 2 * k]) - mx);
                            const float p1 = ex2f(__uint_as_float(sv[16 * c + 2 * k + 1]) - mx);
                            l += p0 + p1;
                            split2(p0, p1, hi[k], lo[k]);
                        }
                        tmem_st8(sb + 8u * c, hi);
                        tmem_st8(plb + 8u * c, lo);
                    }
                    tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bars[B_PFULL + grp]);
#ifdef MS_TRACE
                    if (trace_on) trace[(size_t)t * 16 + 10] += clock64() - soft0;
#endif
                    // o_j = O[group] / l_j (the other group's warps fill the wait for the P V)
                    MS_ACC(9, mbar_wait_bounded(&bars[B_OFULL + grp], ph_og));
                    ph_og ^= 1u;
                    tc_fence_after();
                    uint32_t ov[16], hi[8], lo[8];
                    tmem_ld16(ob, ov);
                    const float inv = __fdividef(1.f, l);
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        split2(__uint_as_float(ov[2 * k]) * inv, __uint_as_float(ov[2 * k + 1]) * inv, hi[k], lo[k]);
                    tmem_st8(tl + C_QH + 8u * j, hi);
                    tmem_st8(tl + C_QL + 8u * j, lo);
                }
                MS_STAMP(3);
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars[B_OREADY]);
                MS_STAMP(4);

                // ---- pointer logits (attention.py:280-282, decoding.py:141-191): this group's NH columns, partial results
                //      exchanged through shared memory, then selection and transition in both threads of the start
                mbar_wait_bounded(&bars[B_LGFULL], ph_lg);
                ph_lg ^= 1u;
                tc_fence_after();
                MS_STAMP(5);
                uint32_t lv[NH];
                tmem_ld_half(tl + C_LG + (uint32_t)(NH * grp), lv);
                // the group's 56 mask bits as a local set: bit n <-> node NH grp + n
                const u64 loc = grp ? ((mk_lo >> NH) | (mk_hi << (64 - NH))) : mk_lo;
                const Bits mkl = make_bits(loc, 0ull);
                int given = -1;
                if (MODE == RL4CO_MODE_EVALUATE && !fin) {
                    const int64_t gv = a.actions_in[(size_t)r * T + t];
                    const bool bad = gv < 0 || gv >= N;
                    status |= bad ? RL4CO_ST_ACTION_RANGE : 0;
                    given = bad ? 0 : (int)gv;
                }
                const int given_loc = given - NH * grp;         // index inside this group's columns (or outside [0, NH))
                // x = clip tanh(z) = clip - 2 clip / (2^(2 log2(e) z) + 1), z = logit / sqrt(E); e = exp(x - clip) (the
                // clip bound is the log-sum-exp shift: no max pass); masked: e = 0.  Strictly branch free: every start
                // has its own mask, a branch per element would diverge on almost every one of them.
                float sum = 0.f, best = -INFINITY, e_a = 0.f, nan_probe = 0.f;
                int act = 0;
#pragma unroll
                for (int n = 0; n < NH; ++n) {
                    const float v = __uint_as_float(lv[n]);
                    nan_probe = fmaf(v, 0.f, nan_probe);         // NaN / inf anywhere in the row -> NaN (pad columns are 0)
                    const float e2 = ex2f(v * (kInvSqrtE * 2.f * kLog2e));
                    const float x = fmaf(-2.f * clip, rcpf(e2 + 1.f), clip);
                    const bool alive = bit_at(mkl, n);
                    const float e = alive ? ex2f((x - clip) * kLog2e) : 0.f;
                    sum += e;
                    if (MODE == RL4CO_MODE_SAMPLING) lv[n] = __float_as_uint(e);
                    if (MODE == RL4CO_MODE_GREEDY) {
                        const float xm = alive ? x : -INFINITY;
                        const bool gt = xm > best;               // strict: lowest index on ties
                        best = gt ? xm : best;
                        act = gt ? n : act;
                        e_a = gt ? e : e_a;
                    } else if (MODE == RL4CO_MODE_EVALUATE) {
                        e_a = (n == given_loc) ? e : e_a;
                    }
                }
                status |= (nan_probe != nan_probe) ? RL4CO_ST_NAN_LOGITS : 0;
                ms.xch[grp][sidx] = make_float4(sum, best, __int_as_float(act + NH * grp), e_a);
                consumers_sync();
                const float4 p0 = ms.xch[0][sidx], p1 = ms.xch[1][sidx];
                sum = p0.x + p1.x;                              // same order in both threads of the start
                if (MODE == RL4CO_MODE_GREEDY) {
                    const bool second = p1.y > p0.y;            // strict: the lower node index wins a tie
                    act = __float_as_int(second ? p1.z : p0.z);
                    e_a = second ? p1.w : p0.w;
                } else if (MODE == RL4CO_MODE_EVALUATE) {
                    act = given < 0 ? 0 : given;
                    e_a = given >= NH ? p1.w : p0.w;
                    status |= (!fin && e_a == 0.f) ? RL4CO_ST_INFEASIBLE_ACTION : 0;
                } else {
                    // Categorical(exp(logp)) by inverse CDF over the start's whole row (decoding.py:383-397): one uniform
                    // per rollout and step; the first feasible node whose running sum of e passes u * sum (the last
                    // feasible one if rounding leaves the running sum short).  Each group scans its own columns (still in
                    // registers) against its share of the target, the winner is agreed through shared memory.
                    const float target = philox_uniform(a.seed, a.offset, (uint64_t)r, (uint32_t)t) * sum - (grp ? p0.x : 0.f);
                    float cum = 0.f;
                    bool found = false;
                    int pick = -1;
                    float e_pick = 0.f;
#pragma unroll
                    for (int n = 0; n < NH; ++n) {
                        const float e = __uint_as_float(lv[n]);
                        cum += e;
                        const bool pos = e > 0.f, take = pos && !found;
                        pick = take ? n : pick;
                        e_pick = take ? e : e_pick;
                        found = found || (pos && cum > target);
                    }
                    consumers_sync();                           // everybody has read the first exchange
                    ms.xch[grp][sidx] = make_float4(found ? 1.f : 0.f, e_pick, __int_as_float(pick < 0 ? -1 : pick + NH * grp), 0.f);
                    consumers_sync();
                    const float4 s0 = ms.xch[0][sidx], s1 = ms.xch[1][sidx];
                    // group 0's columns come first: its hit wins; otherwise group 1's hit or, rounding, its last feasible
                    // node; a group without feasible nodes (pick -1) defers to the other
                    const bool use1 = !(s0.x > 0.f) && __float_as_int(s1.z) >= 0;
                    const bool ok0 = __float_as_int(s0.z) >= 0;
                    act = use1 || !ok0 ? __float_as_int(s1.z) : __float_as_int(s0.z);
                    e_a = use1 || !ok0 ? s1.y : s0.y;
                }
                MS_STAMP(6);
                if (!fin) {
                    const float lp = __logf(e_a) - __logf(sum);
                    if (!(lp > -1000.f)) status |= RL4CO_ST_NEG_INF_LOGP;
                    if (writer) {
                        a.actions[(size_t)r * T + t] = act;
                        if (a.logp) a.logp[(size_t)r * T + t] = lp;
                    }
                    llsum += lp;
                    apply(act);
                    MS_STAMP(7);
                    ++t;
                }
            }

            // ---- outputs of this rollout
            if (writer) {
                const float2 home = CVRP ? ms.locs[0] : ms.locs[first];    // cvrp/env.py:151-160, tsp/env.py:157-163
                len += leg(prev.x, prev.y, home.x, home.y);
                a.reward[r] = -len;
                a.ll[r] = llsum;
                if (a.steps) a.steps[r] = t;
                for (int q = t; q < T; ++q) {                   // pad with depot / log-prob 0 after done
                    a.actions[(size_t)r * T + q] = 0;
                    if (a.logp) a.logp[(size_t)r * T + q] = 0.f;
                }
            }
            if (active && status) raise_status(a.status, status);   // both threads of a start: each saw half of the logits
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

}  // namespace mstc

bool rollout_ms_tc_supports(const rl4co_am_rollout_args &a) {
    if (!a.multistart || a.S < 1 || a.S > 128) return false;
    if (a.start_nodes != nullptr) return false;      // caller-chosen start nodes: the register-file kernel
    if (a.env != RL4CO_ENV_TSP && a.env != RL4CO_ENV_CVRP) return false;
#ifndef MS_TRACE
    if (a.logp_all != nullptr) return false;
#endif
    if (a.N < 2 || a.N > mstc::NK) return false;
    if (!(a.tanh_clipping > 0.f) || a.temperature != 1.f || a.mask_logits == 0) return false;
    return a.n_tables >= (a.env == RL4CO_ENV_TSP ? 5 : 4);
}

int launch_rollout_ms_tc(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    const int64_t grid = a.B < sm_count ? a.B : sm_count;      // persistent: one CTA per SM, instances round robin
    void (*kernel)(const rl4co_am_rollout_args) = nullptr;
    if (a.env == RL4CO_ENV_CVRP)
        kernel = a.mode == RL4CO_MODE_GREEDY     ? mstc::am_rollout_ms_tc_kernel<RL4CO_ENV_CVRP, RL4CO_MODE_GREEDY>
                 : a.mode == RL4CO_MODE_SAMPLING ? mstc::am_rollout_ms_tc_kernel<RL4CO_ENV_CVRP, RL4CO_MODE_SAMPLING>
                                                 : mstc::am_rollout_ms_tc_kernel<RL4CO_ENV_CVRP, RL4CO_MODE_EVALUATE>;
    else
        kernel = a.mode == RL4CO_MODE_GREEDY     ? mstc::am_rollout_ms_tc_kernel<RL4CO_ENV_TSP, RL4CO_MODE_GREEDY>
                 : a.mode == RL4CO_MODE_SAMPLING ? mstc::am_rollout_ms_tc_kernel<RL4CO_ENV_TSP, RL4CO_MODE_SAMPLING>
                                                 : mstc::am_rollout_ms_tc_kernel<RL4CO_ENV_TSP, RL4CO_MODE_EVALUATE>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, mstc::SMEM_BYTES);
    if (e != cudaSuccess) {
        set_error("am_rollout(ms_tc): cannot reserve %d B of shared memory: %s", mstc::SMEM_BYTES, cudaGetErrorString(e));
        return 2;
    }
    kernel<<<(unsigned)grid, mstc::kThreads, mstc::SMEM_BYTES, stream>>>(a);
    RL4CO_CHECK_LAUNCH("am_rollout(ms_tc)");
    return 0;
}

}  // namespace rl4co
