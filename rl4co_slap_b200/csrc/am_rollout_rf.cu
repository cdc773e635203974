// K2 (single start) - persistent Attention-Model rollout with REGISTER-FILE resident tables.
//
// One CTA of 256 threads (8 warps) per SM, one instance at a time, the whole trajectory
// inside the kernel.  The instance's three per-node tables (glimpse K, glimpse V, folded
// logit K: 3 x N x 128 fp32 = 150 KB at N=100) live in the register file for the whole
// trajectory (shared-memory bandwidth would cap a step at 3*N*E*4 B / 128 B/clk = 1200 clk).
//
// Ownership: warp w = attention head w (dims 16w..16w+15); lane = (hh, g) with hh = lane>>4
// the half of the head (8 dims) and g = lane&15 the node group; the thread owns nodes
// g, g+16, ..., g+16(M-1) and, for each of them, the same 8 dims of K, V and L as packed
// fp32 pairs, so that every mat-vec is FFMA2 (fma.rn.f32x2) work on registers - measured on
// B200: ~85 FMA/clk/SM for FFMA2 against ~64 for three-register FFMA (tools/probe_sm.cu).
//
// One decode step = ONE block barrier:
//   every warp, its own head, no block-level communication:
//     q       = (gctx + ctx_first) + C[cur]      C = per-node context table (the one table
//               gathered by row; staged in shared memory by a TMA bulk copy, double buffered)
//     s[n]    = <q_h, K[n]_h> * 0.25*log2(e)     halves combined with one shuffle per node
//     p[n]    = 2^(s[n] - max_n s[n])            masked nodes -> 0      (attention.py:289-306)
//     o_h     = sum_n p[n] V[n]_h                16-lane butterfly, 8 shuffles
//     pl_h[n] = <o_h, L[n]_h> / sum_n p[n]       project_out folded into L -> shared memory
//   barrier (partial logits double buffered, so one barrier per step is enough)
//   every warp, redundantly, lane l = nodes 4l..4l+3: logit = sum_h pl_h / sqrt(E), tanh clip,
//     mask, temperature (utils/decoding.py:141-191); argmax / Philox inverse-CDF sample / given
//     action; env transition and new mask on its own copy of the state bits.
// The chosen log-probability needs a log-sum-exp; its ingredients (max, sum of exponentials -
// each warp contributes the 16 nodes 16w..16w+15 -, chosen logit) are kept per step in shared
// memory and turned into log-probs after the loop, one thread per step, so no logarithm sits
// on the step's critical path.
//
// Reference op order being reproduced (all fp32): SURVEY.md appendix A;
// rl4co/models/zoo/am/decoder.py:134-200, rl4co/models/nn/attention.py:263-309,
// rl4co/utils/decoding.py:141-191,327-397, rl4co/envs/routing/{tsp,cvrp}/env.py,
// rl4co/envs/warehousing/slap/env.py:38-143, rl4co/envs/routing/op/env.py:74-182 (sibling env, this kernel only).
#include "rollout_util.cuh"

namespace rl4co {
namespace rf {

typedef unsigned long long u64;

constexpr int E = RL4CO_EMBED_DIM;   // 128
constexpr int kThreads = 256;
constexpr int kWarps = 8;
constexpr float kScoreScale = 0.25f * 1.4426950408889634f;   // 1/sqrt(16) * log2(e)
constexpr float kInvSqrtE = 0.08838834764831845f;             // 1/sqrt(128)
constexpr float kLog2e = 1.4426950408889634f;

constexpr int PLS = 144;              // row stride of the partial-logit matrix [head][node], = 16 (mod 32)

struct Layout {
    int xs_rows;        // T: every step keeps its logits (log-probs are formed after the loop); 1: per-step sums
    size_t off_c, off_pl, off_xs, off_o, off_ring, off_mw, off_xa, off_act, off_ain, off_lp, off_locs,
        off_dem, off_wc, off_noise, off_part, off_bar, off_dyn, total;
};
__host__ __device__ inline size_t up16(size_t v) { return (v + 15) & ~(size_t)15; }
constexpr size_t kDynBytes = 3 * 128 * 4 + 8 * 128 * 4;     // sdvrp: dynamic-embedding columns + one demand copy per warp
__host__ __device__ inline Layout make_layout(int N, int T, bool has_c, bool has_dyn = false) {
    Layout L;
    // everything except the logits: does a [T][128] ring of them still fit next to it?
    const size_t fixed = (has_c ? (size_t)2 * N * E * 4 : 0) + (size_t)kWarps * PLS * 4 + E * 4 +
                         up16((size_t)T * kWarps * 4) + 5 * up16((size_t)T * 4) + up16((size_t)N * 8) +
                         128 * 4 + 128 * 4 + 16 * 128 * 4 + 64 * 4 + 16 + (has_dyn ? kDynBytes : 0);
    L.xs_rows = (fixed + (size_t)T * 128 * 4 <= (size_t)227 * 1024) ? T : 1;
    size_t o = 0;
    L.off_c = o;    o += has_c ? (size_t)2 * N * E * 4 : 0;     // double-buffered context table
    L.off_pl = o;   o += (size_t)kWarps * PLS * 4;              // partial logits [head][node]
    L.off_xs = o;   o += (size_t)L.xs_rows * 128 * 4;           // processed logits, one row per step (ring only)
    L.off_o = o;    o += E * 4;                                 // glimpse heads (warp private)
    L.off_ring = o; o += up16((size_t)T * kWarps * 4);          // per step, per warp: sum of exponentials
    L.off_mw = o;   o += up16((size_t)T * 4);                   // per step max logit
    L.off_xa = o;   o += up16((size_t)T * 4);                   // per step chosen logit
    L.off_act = o;  o += up16((size_t)T * 4);
    L.off_ain = o;  o += up16((size_t)T * 4);
    L.off_lp = o;   o += up16((size_t)T * 4);
    L.off_locs = o; o += up16((size_t)N * 8);
    L.off_dem = o;  o += 128 * 4;
    L.off_wc = o;   o += 128 * 4;                               // cvrp: remaining-capacity column, pre-scaled
    L.off_noise = o; o += 16 * 128 * 4;                         // sampling: Gumbel noise of 16 steps [16][128]
    L.off_part = o; o += 64 * 4;
    L.off_bar = o;  o += 16;
    L.off_dyn = o;  o += has_dyn ? kDynBytes : 0;               // sdvrp: [3][128] dynamic columns, [8][128] remaining demands
    L.total = o;
    return L;
}

__device__ __forceinline__ float ex2f(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// tanh(x) = 1 - 2 / (2^(2 log2(e) x) + 1): two MUFU ops; absolute error ~2e-7 (the libm tanhf
// costs ~60 cycles of latency on the step's critical path)
__device__ __forceinline__ float fast_tanh(float x) {
    const float t = ex2f(x * 2.8853900817779268f);
    return 1.f - __fdividef(2.f, t + 1.f);
}
__device__ __forceinline__ float rcpf(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float redux_max(float v) {       // all 32 lanes
    float r;
    asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ unsigned redux_min_u32(unsigned v) {
    unsigned r;
    asm volatile("redux.sync.min.u32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(v));
    return r;
}
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {  // SASS: FFMA2
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) {
    u64 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ u64 fadd2(u64 a, u64 b) {
    u64 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ u64 pack2(float lo, float hi) {
    u64 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
    return d;
}
__device__ __forceinline__ float hsum2(u64 v) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    return lo + hi;
}
__device__ __forceinline__ void unpack2(u64 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
// 32 contiguous bytes in one request (SASS: LDG.E.256): a thread's 8 dims of one table row are exactly one sector
__device__ __forceinline__ void ldg_256(const float *p, u64 (&v)[4]) {
    asm volatile("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(p));
}
__device__ __forceinline__ void prefetch_l1(const void *p) {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void *p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// -DRF_TRACE: thread 0 of CTA 0 writes clock64() at eight points of every step of its first
// instance into a.logp_all (reused as a [T][8] int64 buffer by tools/trace_rollout.py)
// -DRF_ABL=<bits>: timing-only ablations for tools/trace_rollout.py (results are wrong): 1 no head max,
// 2 no glimpse butterfly, 4 no half-head combines, 8 no tanh, 16 no argmax reductions
#ifndef RF_ABL
#define RF_ABL 0
#endif
#define RF_SHFL_XOR(bit, v, o) ((RF_ABL & (bit)) ? (v) : __shfl_xor_sync(kFull, (v), (o)))
#ifdef RF_TRACE
#define RF_TRACE_AT(k) do { if (tid == 0 && blockIdx.x == 0 && it == 0) reinterpret_cast<long long *>(a.logp_all)[t * 8 + (k)] = clock64(); } while (0)
#else
#define RF_TRACE_AT(k) do { } while (0)
#endif

template <int ENV, int M>
__global__ void __launch_bounds__(kThreads, 1)
am_rollout_rf_kernel(const rl4co_am_rollout_args a, const int n_parts) {
    constexpr bool HAS_C = ENV != RL4CO_ENV_SLAP;
    constexpr bool VRP = ENV == RL4CO_ENV_CVRP || ENV == RL4CO_ENV_SDVRP;                // capacity state, depot at both ends
    constexpr bool CAPCOL = VRP || ENV == RL4CO_ENV_OP || ENV == RL4CO_ENV_PCTSP;        // context = [h_cur ; scalar state]
    extern __shared__ __align__(128) unsigned char smem[];
    const int N = a.N, T = a.T;
    constexpr bool DYN = ENV == RL4CO_ENV_SDVRP;    // per-step dynamic embedding of the remaining demands
    const Layout L = make_layout(N, T, HAS_C, DYN);
    float *cbuf = reinterpret_cast<float *>(smem + L.off_c);
    float *spl = reinterpret_cast<float *>(smem + L.off_pl);
    float *sxs = reinterpret_cast<float *>(smem + L.off_xs);
    float *so = reinterpret_cast<float *>(smem + L.off_o);
    float *sring = reinterpret_cast<float *>(smem + L.off_ring);
    float *s_mw = reinterpret_cast<float *>(smem + L.off_mw);
    float *s_xa = reinterpret_cast<float *>(smem + L.off_xa);
    int *s_act = reinterpret_cast<int *>(smem + L.off_act);
    int *s_ain = reinterpret_cast<int *>(smem + L.off_ain);
    float *s_lp = reinterpret_cast<float *>(smem + L.off_lp);
    float2 *s_locs = reinterpret_cast<float2 *>(smem + L.off_locs);
    float *s_dem = reinterpret_cast<float *>(smem + L.off_dem);     // demand of node n (0 for depot / pad)
    float *s_wc = reinterpret_cast<float *>(smem + L.off_wc);
    float4 *s_noise = reinterpret_cast<float4 *>(smem + L.off_noise);
    float *spart = reinterpret_cast<float *>(smem + L.off_part);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + L.off_bar); // one per context buffer
    float *s_wdyn = reinterpret_cast<float *>(smem + L.off_dyn);    // sdvrp: w_k, w_v, Wout^T w_l of the dynamic embedding
    float *s_rem = s_wdyn + 3 * 128;                                // sdvrp: remaining demand per node, one copy per warp

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int hh = lane >> 4, g = lane & 15;
    const int dim0 = 16 * warp + 8 * hh;            // first of the 8 dims this thread owns
    const int64_t Mrows = a.B * (int64_t)N;         // rows per table
    const size_t c_bytes = (size_t)N * E * 4;
    const u64 scale2 = pack2(kScoreScale, kScoreScale);
    // the usual decoding configuration gets a branch-free logit path
    const bool xring = L.xs_rows > 1;
    const bool std_logits = a.tanh_clipping > 0.f && a.temperature == 1.f && a.mask_logits != 0;
    const float clip = a.tanh_clipping;
    const bool sampling = a.mode == RL4CO_MODE_SAMPLING;

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_barrier_init();
    }
    if (CAPCOL && tid < E) s_wc[tid] = a.wcap[tid] * kScoreScale;   // context.py:317-331 (cvrp), :385-397 (op)
    if (DYN) for (int q = tid; q < 3 * E; q += kThreads) s_wdyn[q] = a.wdyn[q];
    __syncthreads();
    unsigned par0 = 0u, par1 = 0u;
    auto stage_context = [&](int64_t b, int buf) {  // thread 0: TMA the context table of instance b
        mbar_expect_tx(&bar[buf], (unsigned)c_bytes);
        const char *src = reinterpret_cast<const char *>(a.tables + ((size_t)3 * Mrows + b * (int64_t)N) * E);
        char *dst = reinterpret_cast<char *>(cbuf + (size_t)buf * N * E);
        for (size_t off = 0; off < c_bytes; off += 16384)
            tma_load_1d(dst + off, src + off, (unsigned)(c_bytes - off < 16384 ? c_bytes - off : 16384), &bar[buf]);
    };
    // work item = (instance, range of its starts): with fewer instances than SMs the starts of an instance are
    // spread over several CTAs (n_parts per instance, each loading the instance's tables for itself)
    const int S = a.S > 1 ? a.S : 1;
    const int s_per = (S + n_parts - 1) / n_parts;
    const int64_t W = a.B * (int64_t)n_parts;
    const int t_first = a.multistart ? 1 : 0;       // multistart: step 0 is the forced POMO start node
    if (HAS_C && tid == 0 && (int64_t)blockIdx.x < W) stage_context(blockIdx.x / n_parts, 0);

    int it = 0;
    for (int64_t w = blockIdx.x; w < W; w += gridDim.x, ++it) {
        const int64_t b = w / n_parts;
        const int s_begin = (int)(w % n_parts) * s_per, s_end = s_begin + s_per < S ? s_begin + s_per : S;
        const int buf = it & 1;
        const float *C = cbuf + (size_t)buf * N * E;
        // ------------------------------------------------ registers <- tables (global / L2)
        u64 K2[M][4], V2[M][4], L2[M][4];
        auto load_tables = [&]() {
            const float *kb = a.tables + ((size_t)0 * Mrows + b * (int64_t)N) * E + dim0;
            const float *vb = a.tables + ((size_t)1 * Mrows + b * (int64_t)N) * E + dim0;
            const float *lb = a.tables + ((size_t)2 * Mrows + b * (int64_t)N) * E + dim0;
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const int n = g + 16 * j;
                if (n < N) {
                    ldg_256(kb + (size_t)n * E, K2[j]);
                    ldg_256(vb + (size_t)n * E, V2[j]);
                    ldg_256(lb + (size_t)n * E, L2[j]);
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) K2[j][i] = V2[j][i] = L2[j][i] = 0ull;
                }
            }
        };
        load_tables();
        // constant part of the query, pre-scaled: qbs = scale * (gctx [+ step-0 placeholder])
        u64 qbs2[4];
        auto set_qb = [&](const float *extra) {      // extra: 8 floats at this thread's dims, or null
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = 0.f;
            if (a.gctx) {
                const float4 g0 = __ldg(reinterpret_cast<const float4 *>(a.gctx + b * E + dim0));
                const float4 g1 = __ldg(reinterpret_cast<const float4 *>(a.gctx + b * E + dim0) + 1);
                v[0] = g0.x; v[1] = g0.y; v[2] = g0.z; v[3] = g0.w; v[4] = g1.x; v[5] = g1.y; v[6] = g1.z; v[7] = g1.w;
            }
            if (extra) {
                const float4 e0 = __ldg(reinterpret_cast<const float4 *>(extra));
                const float4 e1 = __ldg(reinterpret_cast<const float4 *>(extra) + 1);
                v[0] += e0.x; v[1] += e0.y; v[2] += e0.z; v[3] += e0.w; v[4] += e1.x; v[5] += e1.y; v[6] += e1.z; v[7] += e1.w;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) qbs2[i] = pack2(v[2 * i] * kScoreScale, v[2 * i + 1] * kScoreScale);
        };
        // ------------------------------------------------ small per-instance state -> smem
        __syncthreads();                            // previous instance's epilogue is done with smem
        if (tid < 128) {
            if (ENV == RL4CO_ENV_OP)    // op: length budget when ARRIVING at node n (op/env.py:129-133), pads never reachable
                s_dem[tid] = tid < N ? a.max_length[b * (int64_t)N + tid] : -INFINITY;
            else if (ENV == RL4CO_ENV_PCTSP)    // pctsp: prize collected at node n (0 for the depot, pctsp/env.py:114-117)
                s_dem[tid] = tid < N ? a.prize[b * (int64_t)N + tid] : 0.f;
            else
                s_dem[tid] = (VRP && tid >= 1 && tid < N) ? a.demand[b * (int64_t)(N - 1) + tid - 1] : 0.f;
        }
        {
            const float2 *lrow = reinterpret_cast<const float2 *>(a.locs + b * (int64_t)N * 2);
            for (int n = tid; n < N; n += kThreads) s_locs[n] = lrow[n];
        }
        // next instance: context table -> the other buffer (its previous user finished before
        // the barrier above), K/V/L rows -> L2
        if (tid == 0 && w + gridDim.x < W) {
            const int64_t nb = (w + gridDim.x) / n_parts;
            if (HAS_C) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                stage_context(nb, buf ^ 1);
            }
            for (int tb = 0; tb < 3; ++tb) {
                const char *src = reinterpret_cast<const char *>(a.tables + ((size_t)tb * Mrows + nb * (int64_t)N) * E);
                for (size_t off = 0; off < c_bytes; off += 32768)
                    prefetch_l2(src + off, (unsigned)(c_bytes - off < 32768 ? c_bytes - off : 32768));
            }
        }
        __syncthreads();

        // ------------------------------------------------ env state, replicated in every warp
        // tsp/env.py:95-120, cvrp/env.py:107-135, slap/env.py:95-129.  Two copies of the mask:
        //   owner layout  - bit j of a word <-> node g + 16 j   (glimpse mask of the mat-vec phase)
        //   reader layout - bit k of a word <-> node 4 lane + k (logit mask of the selection phase)
        unsigned own = 0, rown = 0;
#pragma unroll
        for (int j = 0; j < M; ++j) own |= (g + 16 * j < N) ? (1u << j) : 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k) rown |= (4 * lane + k < N) ? (1u << k) : 0u;
        unsigned alive = own, ralive = rown, vis = 0, rvis = 0;
        // cvrp: vehicle capacity / used capacity; op: the depot's length budget / tour length so far;
        // pctsp: required prize / collected prize
        const float cap = VRP ? a.vehicle_capacity[b]
                        : (ENV == RL4CO_ENV_OP) ? a.max_length[b * (int64_t)N]
                        : (ENV == RL4CO_ENV_PCTSP) ? a.prize_required[b] : 0.f;
        float used = 0.f;
        auto cvrp_masks = [&](int at) {             // cvrp/env.py:137-149, both layouts
            unsigned f = 0, rf = 0;
#pragma unroll
            for (int j = 0; j < M; ++j) {           // demands re-read from shared memory: the register file is full
                const bool exceeds = __fadd_rn(s_dem[g + 16 * j], used) > cap;
                if (!exceeds && !((vis >> j) & 1u)) f |= 1u << j;
            }
            const float4 rdm = *reinterpret_cast<const float4 *>(s_dem + 4 * lane);
            if (!(__fadd_rn(rdm.x, used) > cap) && !(rvis & 1u)) rf |= 1u;
            if (!(__fadd_rn(rdm.y, used) > cap) && !(rvis & 2u)) rf |= 2u;
            if (!(__fadd_rn(rdm.z, used) > cap) && !(rvis & 4u)) rf |= 4u;
            if (!(__fadd_rn(rdm.w, used) > cap) && !(rvis & 8u)) rf |= 8u;
            f &= own;
            rf &= rown;
            if (g == 0) f &= ~1u;
            if (lane == 0) rf &= ~1u;
            const bool depot = !(at == 0 && __any_sync(kFull, rf != 0));
            if (g == 0 && depot) f |= 1u;
            if (lane == 0 && depot) rf |= 1u;
            alive = f;
            ralive = rf;
        };
        auto op_masks = [&](int at) {               // op/env.py:154-170, both layouts
            const float2 pc = s_locs[at];
            const bool closed = (__shfl_sync(kFull, rvis, 0) & 1u) != 0u;   // depot visited: nothing else is feasible
            unsigned f = 0, rf = 0;
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const int n = g + 16 * j;
                const float2 p = s_locs[n < N ? n : 0];
                const bool exceeds = __fadd_rn(used, leg_fma(pc.x, pc.y, p.x, p.y)) > s_dem[n];
                if (!exceeds && !((vis >> j) & 1u)) f |= 1u << j;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int n = 4 * lane + k;
                const float2 p = s_locs[n < N ? n : 0];
                const bool exceeds = __fadd_rn(used, leg_fma(pc.x, pc.y, p.x, p.y)) > s_dem[n];
                if (!exceeds && !((rvis >> k) & 1u)) rf |= 1u << k;
            }
            f &= own;
            rf &= rown;
            if (closed) { f = 0u; rf = 0u; }
            if (g == 0) f |= 1u;                    // the depot can always be visited (:168-169)
            if (lane == 0) rf |= 1u;
            alive = f;
            ralive = rf;
        };
        auto pctsp_masks = [&]() {                  // pctsp/env.py:141-148, both layouts
            const bool closed = (__shfl_sync(kFull, rvis, 0) & 1u) != 0u;   // back at the depot: nothing else is feasible
            const bool all_seen = __all_sync(kFull, (rvis | (lane == 0 ? 1u : 0u)) == rown);
            const bool depot = !(used < 1.0f && !all_seen);      // closed while the prize is short and customers remain
            unsigned f = closed ? 0u : (own & ~vis), rf = closed ? 0u : (rown & ~rvis);
            if (g == 0) f = depot ? (f | 1u) : (f & ~1u);
            if (lane == 0) rf = depot ? (rf | 1u) : (rf & ~1u);
            alive = f;
            ralive = rf;
        };
        float *rem_w = s_rem + warp * 128;           // sdvrp: this warp's copy of demand_with_depot
        auto sdvrp_masks = [&](int at) {            // sdvrp/env.py:132-140, both layouts; done = no demand left (:89)
            const bool full = used >= cap;
            unsigned f = 0, rf = 0;
#pragma unroll
            for (int j = 0; j < M; ++j)
                if (!(rem_w[g + 16 * j] == 0.f) && !full) f |= 1u << j;
            const float4 r4 = *reinterpret_cast<const float4 *>(rem_w + 4 * lane);
            if (!(r4.x == 0.f) && !full) rf |= 1u;
            if (!(r4.y == 0.f) && !full) rf |= 2u;
            if (!(r4.z == 0.f) && !full) rf |= 4u;
            if (!(r4.w == 0.f) && !full) rf |= 8u;
            f &= own;
            rf &= rown;
            if (g == 0) f &= ~1u;
            if (lane == 0) rf &= ~1u;
            const bool depot = !(at == 0 && __any_sync(kFull, rf != 0));
            if (g == 0 && depot) f |= 1u;
            if (lane == 0 && depot) rf |= 1u;
            alive = f;
            ralive = rf;
            return !__any_sync(kFull, r4.x > 0.f || r4.y > 0.f || r4.z > 0.f || r4.w > 0.f);
        };
        // rows of node n with the dynamic embedding of demand d added: X_n = X_stat_n + d w_x (dynamic.py:61-79,
        // am/decoder.py:148-160); product and sum rounded separately like the reference's Linear(1, 3E) + add
        auto dyn_row = [&](u64 (&dst)[4], const u64 (&stat)[4], int tab, float d) {
            const ulonglong2 w0 = *reinterpret_cast<const ulonglong2 *>(s_wdyn + tab * 128 + dim0);
            const ulonglong2 w1 = *(reinterpret_cast<const ulonglong2 *>(s_wdyn + tab * 128 + dim0) + 1);
            const u64 d2 = pack2(d, d);
            dst[0] = fadd2(stat[0], fmul2(d2, w0.x));
            dst[1] = fadd2(stat[1], fmul2(d2, w0.y));
            dst[2] = fadd2(stat[2], fmul2(d2, w1.x));
            dst[3] = fadd2(stat[3], fmul2(d2, w1.y));
        };
        int cur = 0, t = 0, status = 0;
        bool done = false, use_c = CAPCOL;
        const int t_end = (ENV == RL4CO_ENV_TSP) ? (N < T ? N : T) : (ENV == RL4CO_ENV_SLAP ? (a.P < T ? a.P : T) : T);   // cvrp, op: until done

        if (HAS_C) {                                // context table landed?
            if (buf == 0) { mbar_wait(&bar[0], par0); par0 ^= 1u; }
            else          { mbar_wait(&bar[1], par1); par1 ^= 1u; }
        }
        // transition of the chosen (or forced) action, both mask layouts
        auto apply = [&](int act) {
            const unsigned abit = (g == (act & 15)) ? (1u << (act >> 4)) : 0u;
            const unsigned rbit = (lane == (act >> 2)) ? (1u << (act & 3)) : 0u;
            if (ENV == RL4CO_ENV_TSP) {                 // tsp/env.py:67-93
                alive &= ~abit;
                ralive &= ~rbit;
                if (t == 0) {
                    // from step 1 on: ctx = Wc [h_first ; h_cur] (context.py:123-131); the
                    // first-node half is constant for the rest of the trajectory
                    set_qb(a.tables + ((size_t)4 * Mrows + b * (int64_t)N + act) * E + dim0);
                    use_c = true;
                }
            } else if (ENV == RL4CO_ENV_CVRP) {         // cvrp/env.py:73-105,137-149
                const float sel = s_dem[act == 0 ? 1 : act];
                used = __fmul_rn(__fadd_rn(used, sel), act != 0 ? 1.f : 0.f);
                vis |= abit;
                rvis |= rbit;
                cvrp_masks(act);
                done = __all_sync(kFull, rvis == rown);
            } else if (ENV == RL4CO_ENV_SDVRP) {        // sdvrp/env.py:63-101
                const float *srow = a.tables + (b * (int64_t)N + act) * E + dim0;
                if (act != 0 && g == (act & 15)) {          // the rows rebuilt below: start them towards L1 now (no registers)
                    prefetch_l1(srow);
                    prefetch_l1(srow + (size_t)Mrows * E);
                    prefetch_l1(srow + (size_t)2 * Mrows * E);
                }
                const float sel = rem_w[act];               // the depot slot is 0
                const float delivered = fminf(sel, __fsub_rn(cap, used));
                used = __fmul_rn(__fadd_rn(used, delivered), act != 0 ? 1.f : 0.f);
                const float left = __fadd_rn(sel, -delivered);
                __syncwarp();
                if (lane == 0) rem_w[act] = left;
                __syncwarp();
                done = sdvrp_masks(act);
                // the visited node's keys / values / logit keys follow its remaining demand: its owner threads rebuild
                // their 8 dims from the static rows (tables 0-2, L2 resident) - one row per step, nothing else changes
                if (!done && act != 0 && g == (act & 15)) {
                    u64 sk[4], sv[4], sl[4];
                    ldg_256(srow, sk);
                    ldg_256(srow + (size_t)Mrows * E, sv);
                    ldg_256(srow + (size_t)2 * Mrows * E, sl);
                    dyn_row(sk, sk, 0, left);
                    dyn_row(sv, sv, 1, left);
                    dyn_row(sl, sl, 2, left);
                    const int row = act >> 4;
#pragma unroll
                    for (int j = 0; j < M; ++j) {       // selects, not indexing: the tables must stay in registers
                        const bool hit = j == row;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            K2[j][i] = hit ? sk[i] : K2[j][i];
                            V2[j][i] = hit ? sv[i] : V2[j][i];
                            L2[j][i] = hit ? sl[i] : L2[j][i];
                        }
                    }
                }
            } else if (ENV == RL4CO_ENV_OP) {           // op/env.py:74-108
                const float2 pp = s_locs[cur], pn = s_locs[act];
                used = __fadd_rn(used, leg_fma(pp.x, pp.y, pn.x, pn.y));
                vis |= abit;
                rvis |= rbit;
                done = act == 0 && t > 0;               // back at the depot, except on the very first step
                if (!done) op_masks(act);
            } else if (ENV == RL4CO_ENV_PCTSP) {        // pctsp/env.py:69-101
                used = __fadd_rn(used, s_dem[act]);
                vis |= abit;
                rvis |= rbit;
                done = act == 0 && t > 0;
                if (!done) pctsp_masks();
            } else {                                    // slap/env.py:38-93
                alive &= ~abit;
                ralive &= ~rbit;
            }
            cur = act;
        };

        // log-prob ingredients of step tp (its logits px, their max pgm, its action pact): this warp's
        // share of the sum of exponentials = nodes 16w..16w+15 = lanes 4w..4w+3, chosen logit, action
        float px[4] = {0.f, 0.f, 0.f, 0.f}, pgm = 0.f;
        int pact = 0;
        auto book = [&](int tp) {
            float e[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) e[k] = (px[k] == -INFINITY) ? 0.f : ex2f((px[k] - pgm) * kLog2e);
            float esum = (e[0] + e[1]) + (e[2] + e[3]);
            esum += __shfl_xor_sync(kFull, esum, 1);
            esum += __shfl_xor_sync(kFull, esum, 2);
            if (lane == 4 * warp) sring[tp * kWarps + warp] = esum;
            if (warp == 0) {
                float xa = pgm;
                if (a.mode != RL4CO_MODE_GREEDY) {
                    const int ak = pact & 3;
                    const float mine = ak == 0 ? px[0] : ak == 1 ? px[1] : ak == 2 ? px[2] : px[3];
                    xa = __shfl_sync(kFull, mine, pact >> 2);
                }
                if (lane == 0) { s_act[tp] = pact; s_xa[tp] = xa; s_mw[tp] = pgm; }
            }
        };

        for (int s = s_begin; s < s_end; ++s) {
        const int64_t r = (int64_t)s * a.B + b;     // flat rollout index, start-major (utils/ops.py:11-34)
        auto gen_noise = [&](int step) {
            float Gn[4];
            gumbel4(a.seed, a.offset, (uint64_t)r, (uint32_t)step, lane, Gn);
            s_noise[(step & 15) * 32 + lane] = make_float4(Gn[0], Gn[1], Gn[2], Gn[3]);
        };
        // ------------------------------------------------ reset of the per-rollout state
        if (s != s_begin) __syncthreads();          // the previous rollout's epilogue is done with shared memory
        if (sampling) {                             // Gumbel noise of steps 0..15 (warp w: steps w and w + 8)
            gen_noise(warp);
            gen_noise(warp + 8);
            __syncthreads();
        }
        if (a.mode == RL4CO_MODE_EVALUATE) {
            for (int q = tid; q < T; q += kThreads) {
                int64_t gv = a.actions_in[(size_t)r * T + q];
                s_ain[q] = (gv < 0 || gv >= N) ? -1 : (int)gv;
            }
            __syncthreads();
        }
        alive = own; ralive = rown; vis = 0; rvis = 0; used = 0.f;
        if (ENV == RL4CO_ENV_SLAP) {                // slap: depot never assignable
            if (g == 0) alive &= ~1u;
            if (lane == 0) ralive &= ~1u;
        }
        if (ENV == RL4CO_ENV_CVRP) cvrp_masks(0);
        if (ENV == RL4CO_ENV_SDVRP) {               // remaining demands = demands, every row gets its dynamic term
            // several starts per instance (am/decoder.py:176-187: keys and values are NOT shared across starts, the
            // episodes modify them): the previous start left its own dynamic terms in the registers - static rows again
            if (s != s_begin) load_tables();
            *reinterpret_cast<float4 *>(rem_w + 4 * lane) = *reinterpret_cast<const float4 *>(s_dem + 4 * lane);
            __syncwarp();
            sdvrp_masks(0);
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const float d = rem_w[g + 16 * j];
                dyn_row(K2[j], K2[j], 0, d);
                dyn_row(V2[j], V2[j], 1, d);
                dyn_row(L2[j], L2[j], 2, d);
            }
        }
        if (ENV == RL4CO_ENV_OP) op_masks(0);
        if (ENV == RL4CO_ENV_PCTSP) pctsp_masks();
        cur = 0; t = 0; status = 0; done = false; use_c = CAPCOL;
        set_qb(ENV == RL4CO_ENV_TSP ? a.q0 + dim0 : nullptr);   // tsp step 0: context.py:111-121
        if (a.multistart) {                         // forced POMO start node, log-prob 0 (utils/decoding.py:287-305)
            const int a0 = a.start_nodes ? (int)a.start_nodes[r] : s % a.num_loc + (ENV == RL4CO_ENV_TSP ? 0 : 1);
            if (tid == 0) { s_act[0] = a0; s_lp[0] = 0.f; }
            apply(a0);
            t = 1;
        }

        // ================================================== decode loop ==================
        while (!done && t < t_end) {
            RF_TRACE_AT(0);
            // sampling = arg-max of logits + Gumbel noise (rollout_util.cuh).  Every warp needs the same 128 values per
            // step: warp w draws the noise of step t0 + 8 + w once every eight steps into a ring of 16 steps; a row
            // is rewritten only after the barriers of the step that read it
            float G[4] = {0.f, 0.f, 0.f, 0.f};
            if (sampling) {
                if ((t & 7) == 0 && t > 0) gen_noise(t + 8 + warp);
                const float4 g4 = s_noise[(t & 15) * 32 + lane];
                G[0] = g4.x; G[1] = g4.y; G[2] = g4.z; G[3] = g4.w;
            }
            // ---- query of this half head (decoder.py:141-142), pre-scaled
            u64 q2[4];
            if (HAS_C && use_c) {
                const ulonglong2 c0 = *reinterpret_cast<const ulonglong2 *>(C + (size_t)cur * E + dim0);
                const ulonglong2 c1 = *(reinterpret_cast<const ulonglong2 *>(C + (size_t)cur * E + dim0) + 1);
                q2[0] = ffma2(c0.x, scale2, qbs2[0]);
                q2[1] = ffma2(c0.y, scale2, qbs2[1]);
                q2[2] = ffma2(c1.x, scale2, qbs2[2]);
                q2[3] = ffma2(c1.y, scale2, qbs2[3]);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) q2[i] = qbs2[i];
            }
            if (CAPCOL) {                               // + (capacity - used) * remaining-capacity column (op: length)
                // cvrp: capacity - used; op: length budget - tour length; pctsp: max(required - collected prize, 0)
                const float rem = (ENV == RL4CO_ENV_PCTSP) ? fmaxf(__fsub_rn(cap, used), 0.f) : __fsub_rn(cap, used);
                const u64 rem2 = pack2(rem, rem);
                const ulonglong2 w0 = *reinterpret_cast<const ulonglong2 *>(s_wc + dim0);
                const ulonglong2 w1 = *(reinterpret_cast<const ulonglong2 *>(s_wc + dim0) + 1);
                q2[0] = ffma2(w0.x, rem2, q2[0]);
                q2[1] = ffma2(w0.y, rem2, q2[1]);
                q2[2] = ffma2(w1.x, rem2, q2[2]);
                q2[3] = ffma2(w1.y, rem2, q2[3]);
            }
            // ---- glimpse scores of the own nodes (attention.py:289-306), base-2 scaled
            float s[M];
            {
                u64 acc[M];
#pragma unroll
                for (int j = 0; j < M; ++j) acc[j] = 0ull;
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < M; ++j) acc[j] = ffma2(K2[j][i], q2[i], acc[j]);
                if (!xring && t > t_first) book(t - 1); // overlaps with the FFMA2 latency above
#pragma unroll
                for (int j = 0; j < M; ++j) s[j] = hsum2(acc[j]);
#pragma unroll
                for (int j = 0; j < M; ++j) s[j] += RF_SHFL_XOR(4, s[j], 16);
            }
            float mx = -INFINITY;
#pragma unroll
            for (int j = 0; j < M; ++j) {
                s[j] = ((alive >> j) & 1u) ? s[j] : -INFINITY;
                mx = fmaxf(mx, s[j]);
            }
            if (!(RF_ABL & 1)) mx = redux_max(mx);
            RF_TRACE_AT(1);
            float psum = 0.f;
            u64 oacc[4] = {0ull, 0ull, 0ull, 0ull};
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const float p = ex2f(s[j] - mx);        // masked: 2^-inf = 0
                psum += p;
                const u64 pp = pack2(p, p);
#pragma unroll
                for (int i = 0; i < 4; ++i) oacc[i] = ffma2(V2[j][i], pp, oacc[i]);
            }
            // ---- reduce the 8 head dims (and the normaliser) over the 16 node groups
            {
                float v[8];
#pragma unroll
                for (int i = 0; i < 4; ++i) unpack2(oacc[i], v[2 * i], v[2 * i + 1]);
#pragma unroll
                for (int o = 8; o > 0; o >>= 1) psum += __shfl_xor_sync(kFull, psum, o);
                const bool b3 = g & 8, b2 = g & 4, b1 = g & 2;
                float w4[4], w2[2], w1;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float send = b3 ? v[k] : v[k + 4], keep = b3 ? v[k + 4] : v[k];
                    w4[k] = keep + RF_SHFL_XOR(2, send, 8);
                }
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const float send = b2 ? w4[k] : w4[k + 2], keep = b2 ? w4[k + 2] : w4[k];
                    w2[k] = keep + RF_SHFL_XOR(2, send, 4);
                }
                {
                    const float send = b1 ? w2[0] : w2[1], keep = b1 ? w2[1] : w2[0];
                    w1 = keep + RF_SHFL_XOR(2, send, 2);
                }
                w1 += RF_SHFL_XOR(2, w1, 1);
                // this lane now holds head dim (b3 ? 4 : 0) + (b2 ? 2 : 0) + (b1 ? 1 : 0) of its half,
                // un-normalised; 1/psum is applied to the partial logits
                so[dim0 + (b3 ? 4 : 0) + (b2 ? 2 : 0) + (b1 ? 1 : 0)] = w1;
            }
            __syncwarp();
            RF_TRACE_AT(2);
            // ---- partial pointer logits of this head, project_out folded (attention.py:276-282)
            {
                const ulonglong2 o0 = *reinterpret_cast<const ulonglong2 *>(so + dim0);
                const ulonglong2 o1 = *(reinterpret_cast<const ulonglong2 *>(so + dim0) + 1);
                const u64 o2[4] = {o0.x, o0.y, o1.x, o1.y};
                u64 acc[M];
#pragma unroll
                for (int j = 0; j < M; ++j) acc[j] = 0ull;
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < M; ++j) acc[j] = ffma2(L2[j][i], o2[i], acc[j]);
                const float inv = __fdividef(1.f, psum);    // softmax normaliser of this head
                float pl[M];
#pragma unroll
                for (int j = 0; j < M; ++j) pl[j] = hsum2(acc[j]) * inv;
#pragma unroll
                for (int j = 0; j < M; ++j) pl[j] += RF_SHFL_XOR(4, pl[j], 16);
                float *row = spl + warp * PLS + g;
                if (hh == 0) {
#pragma unroll
                    for (int j = 0; j < M; ++j) row[16 * j] = pl[j];
                }
            }
            RF_TRACE_AT(3);
            __syncthreads();
            RF_TRACE_AT(4);

            float *sxrow = sxs + (xring ? t * 128 : 0);
            // ---- warps 0-3, thread = node 32 warp + lane (= g + 16 (2 warp + hh), so its feasibility bit is in the
            //      owner-layout mask): logit = sum_h pl_h / sqrt(E), then process_logits (utils/decoding.py:141-191)
            if (warp < 4) {
                const float *col = spl + tid;
                const float sum = ((col[0] + col[2 * PLS]) + (col[4 * PLS] + col[6 * PLS])) +
                                  ((col[PLS] + col[3 * PLS]) + (col[5 * PLS] + col[7 * PLS]));
                const int jb = 2 * warp + hh;
                const bool exists = (own >> jb) & 1u, feasible = (alive >> jb) & 1u;
                float xv;
                if (std_logits) {
                    // clip * tanh(z) = clip - 2 clip / (2^(2 log2(e) z) + 1), z = sum / sqrt(E)
                    const float e2 = ex2f(sum * (kInvSqrtE * 2.f * kLog2e));
                    xv = fmaf(-2.f * clip, rcpf(e2 + 1.f), clip);
                    if (RF_ABL & 8) xv = sum;
                    if (exists && sum != sum) status |= RL4CO_ST_NAN_LOGITS;
                    xv = feasible ? xv : -INFINITY;
                } else {
                    xv = sum * kInvSqrtE;
                    if (exists && xv != xv) status |= RL4CO_ST_NAN_LOGITS;
                    if (clip > 0.f) xv = fast_tanh(xv) * clip;
                    xv = __fdiv_rn(xv, a.temperature);
                    xv = (exists && (feasible || !a.mask_logits)) ? xv : -INFINITY;
                }
                sxrow[tid] = xv;
            }
            RF_TRACE_AT(5);
            __syncthreads();
            RF_TRACE_AT(6);
            // ---- every warp: all logits, lane = nodes 4 lane .. 4 lane + 3
            float x[4];
            {
                const float4 xv = *reinterpret_cast<const float4 *>(sxrow + 4 * lane);
                x[0] = xv.x; x[1] = xv.y; x[2] = xv.z; x[3] = xv.w;
            }
            // the log-sum-exp shift: the max logit; sampling with clipped logits uses their bound instead (no reduction)
            const float gm = (RF_ABL & 16) ? x[0] : (sampling && std_logits) ? clip
                                                    : redux_max(fmaxf(fmaxf(x[0], x[1]), fmaxf(x[2], x[3])));
            // ---- selection (utils/decoding.py:371-397,410-499)
            int act;
            if (RF_ABL & 16) {
                act = t + (x[0] > 1e30f ? 1 : 0);
            } else if (a.mode == RL4CO_MODE_GREEDY) {
                // argmax, lowest index on ties (the log-softmax shift is common to all nodes)
                unsigned cand = 0x7fffffffu;
                cand = x[3] == gm ? 4u * lane + 3u : cand;
                cand = x[2] == gm ? 4u * lane + 2u : cand;
                cand = x[1] == gm ? 4u * lane + 1u : cand;
                cand = x[0] == gm ? 4u * lane : cand;
                const unsigned win = redux_min_u32(cand);
                act = win < (unsigned)N ? (int)win : 0;
            } else if (a.mode == RL4CO_MODE_EVALUATE) {
                act = s_ain[t];
                if (act < 0) { status |= RL4CO_ST_ACTION_RANGE; act = 0; }
            } else {
                // Gumbel-max sample of Categorical(exp(logp)) (rollout_util.cuh): arg-max of logits + noise; the
                // noise was drawn at the top of the step, every warp has the same, so all of them agree
                const float y0 = x[0] + G[0], y1 = x[1] + G[1], y2 = x[2] + G[2], y3 = x[3] + G[3];
                const float ym = redux_max(fmaxf(fmaxf(y0, y1), fmaxf(y2, y3)));
                unsigned cand = 0x7fffffffu;
                cand = y3 == ym ? 4u * lane + 3u : cand;
                cand = y2 == ym ? 4u * lane + 2u : cand;
                cand = y1 == ym ? 4u * lane + 1u : cand;
                cand = y0 == ym ? 4u * lane : cand;
                const unsigned win = redux_min_u32(cand);
                act = win < (unsigned)N ? (int)win : 0;
            }
            // feasibility of the chosen action (utils/decoding.py:377-379,393-395): its owner lane checks
            if (lane == (act >> 2) && !((ralive >> (act & 3)) & 1u)) status |= RL4CO_ST_INFEASIBLE_ACTION;
            // the step's log-prob ingredients are recorded during the next step's mat-vecs (book())
#pragma unroll
            for (int k = 0; k < 4; ++k) px[k] = x[k];
            pgm = gm;
            pact = act;
            if (xring && tid == 0) { s_act[t] = act; s_mw[t] = gm; }
            apply(act);
            RF_TRACE_AT(7);
            ++t;
        }

        // ================================================== epilogue =====================
        if (!xring && t > t_first) book(t - 1);
        __syncthreads();
        const int nact = t;
        // chosen log-probs: logp = (x_a - m) - log(sum exp(x - m))   (utils/decoding.py:191,364-368)
        if (xring) {
            // warp w takes steps w, w + 8, ... in groups of 16: the sums of a group are reduced together and lane u
            // finishes step w + 8 u, so neither the shuffles nor the logarithm are serialised over the steps;
            // exponentials through ex2.approx (relative error 2^-22; masked logits: 2^-inf = 0)
            for (int qb = warp; qb < nact; qb += 16 * kWarps) {
                float es[16], mm[16];
#pragma unroll
                for (int u = 0; u < 16; ++u) {
                    const int q = qb + u * kWarps;
                    es[u] = 0.f;
                    mm[u] = 0.f;
                    if (q >= t_first && q < nact) {
                        const float4 xv = *reinterpret_cast<const float4 *>(sxs + q * 128 + 4 * lane);
                        mm[u] = s_mw[q];
                        es[u] = (ex2f((xv.x - mm[u]) * kLog2e) + ex2f((xv.y - mm[u]) * kLog2e)) +
                                (ex2f((xv.z - mm[u]) * kLog2e) + ex2f((xv.w - mm[u]) * kLog2e));
                    }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1)
#pragma unroll
                    for (int u = 0; u < 16; ++u) es[u] += __shfl_xor_sync(kFull, es[u], o);
                float my_es = 1.f, my_m = 0.f;
#pragma unroll
                for (int u = 0; u < 16; ++u)
                    if (lane == u) { my_es = es[u]; my_m = mm[u]; }
                const int q = qb + lane * kWarps;
                if (lane < 16 && q >= t_first && q < nact) {
                    const float lpa = (sxs[q * 128 + s_act[q]] - my_m) - logf(my_es);
                    s_lp[q] = lpa;
                    if (!(lpa > -1000.f)) status |= RL4CO_ST_NEG_INF_LOGP;
                }
            }
        } else {
            for (int q = tid + t_first; q < nact; q += kThreads) {
                const float4 r0 = *reinterpret_cast<const float4 *>(sring + q * kWarps);
                const float4 r1 = *reinterpret_cast<const float4 *>(sring + q * kWarps + 4);
                const float tot = ((r0.x + r0.y) + (r0.z + r0.w)) + ((r1.x + r1.y) + (r1.z + r1.w));
                const float lpa = (s_xa[q] - s_mw[q]) - logf(tot);
                s_lp[q] = lpa;
                if (!(lpa > -1000.f)) status |= RL4CO_ST_NEG_INF_LOGP;
            }
        }
        float part = 0.f;
        if (ENV == RL4CO_ENV_TSP) {              // tsp/env.py:157-163
            for (int q = tid; q < nact; q += kThreads) {
                float2 p = s_locs[s_act[q]], w = s_locs[s_act[q + 1 == nact ? 0 : q + 1]];
                part += leg(p.x, p.y, w.x, w.y);
            }
        } else if (VRP) {                        // cvrp/env.py:151-160 (depot at both ends; sdvrp inherits it)
            for (int q = tid; q <= nact; q += kThreads) {
                float2 p = q == 0 ? s_locs[0] : s_locs[s_act[q - 1]];
                float2 w = q == nact ? s_locs[0] : s_locs[s_act[q]];
                part += leg(p.x, p.y, w.x, w.y);
            }
        } else if (ENV == RL4CO_ENV_PCTSP) {     // pctsp/env.py:150-174: length + all penalties - saved penalties
            const float *pen = a.penalty + b * (int64_t)N;
            for (int q = tid; q <= nact; q += kThreads) {
                float2 p = q == 0 ? s_locs[0] : s_locs[s_act[q - 1]];
                float2 w = q == nact ? s_locs[0] : s_locs[s_act[q]];
                part += leg(p.x, p.y, w.x, w.y);
                if (q < nact) part -= pen[s_act[q]];
            }
            for (int n = tid + 1; n < N; n += kThreads) part += pen[n];
        } else if (ENV == RL4CO_ENV_OP) {        // op/env.py:172-182: minus the collected prize (negated below)
            for (int q = tid; q < nact; q += kThreads) part -= a.prize[b * (int64_t)N + s_act[q]];
        } else {                                 // slap/env.py:131-143 (assignment[t] = s_act[t])
            const int64_t *pl = a.picklist + b * (int64_t)a.O * a.K;
            for (int o = tid; o < a.O; o += kThreads)
                part += slap_order_len(s_locs, s_act, pl + (int64_t)o * a.K, a.K);
        }
        part = warp_sum(part);
        if (lane == 0) spart[warp] = part;
        status = __reduce_or_sync(kFull, status);
        if (lane == 0) spart[40 + warp] = __int_as_float(status);
        __syncthreads();                         // s_lp complete, partials visible
        if (warp == 0) {
            float llp = 0.f;
            for (int q = lane; q < nact; q += 32) llp += s_lp[q];
            llp = warp_sum(llp);
            if (lane == 0) {
                float len = 0.f;
                int st = 0;
                for (int w = 0; w < kWarps; ++w) { len += spart[w]; st |= __float_as_int(spart[40 + w]); }
                a.reward[r] = -len;
                a.ll[r] = llp;
                if (a.steps) a.steps[r] = nact;
                if (st) raise_status(a.status, st);
            }
        }
        for (int q = tid; q < T; q += kThreads) {    // pad with depot / log-prob 0 after done
            a.actions[(size_t)r * T + q] = q < nact ? (int64_t)s_act[q] : 0;
            if (a.logp) a.logp[(size_t)r * T + q] = q < nact ? s_lp[q] : 0.f;
        }
        if (ENV == RL4CO_ENV_SLAP && a.assignment)
            for (int q = tid; q < a.P; q += kThreads) a.assignment[(size_t)r * a.P + q] = q < nact ? s_act[q] : -1;
        }   // starts of this work item
    }
}

template <int ENV, int M>
static int launch(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    const Layout L = make_layout(a.N, a.T, ENV != RL4CO_ENV_SLAP, ENV == RL4CO_ENV_SDVRP);
    if (L.total > (size_t)227 * 1024) {
        set_error("am_rollout(rf): N=%d T=%d needs %zu B shared memory", a.N, a.T, L.total);
        return 1;
    }
    cudaError_t e = cudaFuncSetAttribute(am_rollout_rf_kernel<ENV, M>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
    if (e != cudaSuccess) {
        set_error("am_rollout(rf): cannot reserve %zu B of shared memory: %s", L.total, cudaGetErrorString(e));
        return 2;
    }
    // fewer instances than ~16 per SM: spread the starts of an instance over several work items
    const int S = a.S > 1 ? a.S : 1;
    int64_t n_parts = 1;
    if (S > 1 && a.B < (int64_t)16 * sm_count) {
        n_parts = ((int64_t)16 * sm_count + a.B - 1) / a.B;
        if (n_parts > S) n_parts = S;
        const int64_t per = (S + n_parts - 1) / n_parts;
        n_parts = (S + per - 1) / per;                       // no empty parts
    }
    const int64_t W = a.B * n_parts;
    const int64_t grid = W < sm_count ? W : sm_count;        // persistent: one CTA per SM
    am_rollout_rf_kernel<ENV, M><<<(unsigned)grid, kThreads, L.total, stream>>>(a, (int)n_parts);
    RL4CO_CHECK_LAUNCH("am_rollout(rf)");
    return 0;
}

__host__ inline int nodes_per_thread(int N) { return N <= 32 ? 2 : N <= 64 ? 4 : N <= 112 ? 7 : 8; }

template <int ENV>
static int launch_env(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    switch (nodes_per_thread(a.N)) {
        case 2: return launch<ENV, 2>(a, stream);
        case 4: return launch<ENV, 4>(a, stream);
        case 7: return launch<ENV, 7>(a, stream);
        default: return launch<ENV, 8>(a, stream);
    }
}

}  // namespace rf

bool rollout_rf_supports(const rl4co_am_rollout_args &a) {
#ifndef RF_TRACE
    if (a.logp_all != nullptr) return false;
#endif
    if (a.N > 128) return false;
    if (a.w_first != nullptr) return false;     // tsp without table 4: the compaction kernel or the shared-memory kernel
    if (a.multistart && a.env == RL4CO_ENV_SLAP) return false;
    return rf::make_layout(a.N, a.T, a.env != RL4CO_ENV_SLAP, a.env == RL4CO_ENV_SDVRP).total <= (size_t)227 * 1024;
}

int launch_rollout_rf(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    switch (a.env) {
        case RL4CO_ENV_TSP: return rf::launch_env<RL4CO_ENV_TSP>(a, stream);
        case RL4CO_ENV_CVRP: return rf::launch_env<RL4CO_ENV_CVRP>(a, stream);
        case RL4CO_ENV_OP: return rf::launch_env<RL4CO_ENV_OP>(a, stream);
        case RL4CO_ENV_PCTSP: return rf::launch_env<RL4CO_ENV_PCTSP>(a, stream);
        case RL4CO_ENV_SDVRP: return rf::launch_env<RL4CO_ENV_SDVRP>(a, stream);
        default: return rf::launch_env<RL4CO_ENV_SLAP>(a, stream);
    }
}

}  // namespace rl4co
