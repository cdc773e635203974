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
// One decode step = 2 block barriers:
//   every warp, its own head, no block-level communication:
//     q       = (gctx + ctx_first) + C[cur]      C = per-node context table (the one table
//               gathered by row; staged in shared memory by a TMA bulk copy, double buffered)
//     s[n]    = <q_h, K[n]_h> * 0.25*log2(e)     halves combined with one shuffle per node
//     p[n]    = 2^(s[n] - max_n s[n])            masked nodes -> 0      (attention.py:289-306)
//     o_h     = sum_n p[n] V[n]_h / sum_n p[n]   16-lane butterfly, 8 shuffles
//     pl_h[n] = <o_h, L[n]_h>                    project_out folded into L -> shared memory
//   barrier 1
//   warp w finishes nodes 16w..16w+15: logit = sum_h pl_h / sqrt(E), tanh clip, mask,
//     temperature (utils/decoding.py:141-191); warp max / argmax / sum of exponentials
//   barrier 2
//   every warp (8 candidates): global argmax / Philox inverse-CDF sample / given action, then
//     the env transition and the new mask on its own copy of the state bits.
// The chosen log-probability needs a log-sum-exp; its ingredients (per-warp max and sum,
// chosen logit) are kept per step in shared memory and turned into log-probs after the loop,
// one thread per step, so no transcendental sits on the step's critical path.
//
// Reference op order being reproduced (all fp32): SURVEY.md appendix A;
// rl4co/models/zoo/am/decoder.py:134-200, rl4co/models/nn/attention.py:263-309,
// rl4co/utils/decoding.py:141-191,327-397, rl4co/envs/routing/{tsp,cvrp}/env.py,
// rl4co/envs/warehousing/slap/env.py:38-143.
#include "rollout_util.cuh"

namespace rl4co {
namespace rf {

typedef unsigned long long u64;

constexpr int E = RL4CO_EMBED_DIM;   // 128
constexpr int kThreads = 256;
constexpr int kWarps = 8;
constexpr float kScoreScale = 0.25f * 1.4426950408889634f;   // 1/sqrt(16) * log2(e)
constexpr float kInvSqrtE = 0.08838834764831845f;             // 1/sqrt(128)

// row stride of the partial-logit matrix: = 16 (mod 32) so the two lane halves hit disjoint banks
__host__ __device__ constexpr int pl_stride(int M) { return 16 * M + ((M & 1) ? 0 : 16); }

struct Layout {
    size_t off_c, off_pl, off_o, off_xs, off_cand, off_pref, off_ring, off_xa, off_act, off_ain,
        off_lp, off_locs, off_dem, off_part, off_bar, total;
};
__host__ __device__ inline size_t up16(size_t v) { return (v + 15) & ~(size_t)15; }
__host__ __device__ inline Layout make_layout(int N, int T, int M, bool has_c) {
    Layout L;
    size_t o = 0;
    L.off_c = o;    o += has_c ? (size_t)2 * N * E * 4 : 0;     // double-buffered context table
    L.off_pl = o;   o += (size_t)16 * pl_stride(M) * 4;         // [2*head + half][node]
    L.off_o = o;    o += E * 4;                                 // glimpse heads (warp private)
    L.off_xs = o;   o += (size_t)16 * M * 4;                    // processed logits of the step
    L.off_cand = o; o += kWarps * 8;                            // (max, argmax | sum) per warp
    L.off_pref = o; o += kWarps * 16 * 4;                       // sampling: prefix sums per warp
    L.off_ring = o; o += up16((size_t)T * kWarps * 8);          // per step, per warp (max, sum)
    L.off_xa = o;   o += up16((size_t)T * 4);                   // per step chosen logit
    L.off_act = o;  o += up16((size_t)T * 4);
    L.off_ain = o;  o += up16((size_t)T * 4);
    L.off_lp = o;   o += up16((size_t)T * 4);
    L.off_locs = o; o += up16((size_t)N * 8);
    L.off_dem = o;  o += up16((size_t)16 * M * 4);
    L.off_part = o; o += 64 * 4;
    L.off_bar = o;  o += 16;
    L.total = o;
    return L;
}

__device__ __forceinline__ float ex2f(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float redux_max(float v) {       // all 32 lanes
    float r;
    asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {  // SASS: FFMA2
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
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
__device__ __forceinline__ void prefetch_l2(const void *p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

template <int ENV, int M>
__global__ void __launch_bounds__(kThreads, 1)
am_rollout_rf_kernel(const rl4co_am_rollout_args a) {
    constexpr int NS = 16 * M;                      // node slots
    constexpr int PLS = pl_stride(M);
    constexpr bool HAS_C = ENV != RL4CO_ENV_SLAP;
    extern __shared__ __align__(128) unsigned char smem[];
    const int N = a.N, T = a.T;
    const Layout L = make_layout(N, T, M, HAS_C);
    float *cbuf = reinterpret_cast<float *>(smem + L.off_c);
    float *spl = reinterpret_cast<float *>(smem + L.off_pl);
    float *so = reinterpret_cast<float *>(smem + L.off_o);
    float *sxs = reinterpret_cast<float *>(smem + L.off_xs);
    float2 *scand = reinterpret_cast<float2 *>(smem + L.off_cand);
    float *spref = reinterpret_cast<float *>(smem + L.off_pref);
    float2 *sring = reinterpret_cast<float2 *>(smem + L.off_ring);
    float *s_xa = reinterpret_cast<float *>(smem + L.off_xa);
    int *s_act = reinterpret_cast<int *>(smem + L.off_act);
    int *s_ain = reinterpret_cast<int *>(smem + L.off_ain);
    float *s_lp = reinterpret_cast<float *>(smem + L.off_lp);
    float2 *s_locs = reinterpret_cast<float2 *>(smem + L.off_locs);
    float *s_dem = reinterpret_cast<float *>(smem + L.off_dem);     // demand of node n (0 for depot / pad)
    float *spart = reinterpret_cast<float *>(smem + L.off_part);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + L.off_bar); // one per context buffer

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int hh = lane >> 4, g = lane & 15;
    const int dim0 = 16 * warp + 8 * hh;            // first of the 8 dims this thread owns
    const int64_t Mrows = a.B * (int64_t)N;         // rows per table
    const size_t c_bytes = (size_t)N * E * 4;
    const u64 scale2 = pack2(kScoreScale, kScoreScale);
    const bool sampling = a.mode == RL4CO_MODE_SAMPLING;

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_barrier_init();
    }
    __syncthreads();
    unsigned parity[2] = {0u, 0u};
    auto stage_context = [&](int64_t b, int buf) {  // thread 0: TMA the context table of instance b
        mbar_expect_tx(&bar[buf], (unsigned)c_bytes);
        const char *src = reinterpret_cast<const char *>(a.tables + ((size_t)3 * Mrows + b * (int64_t)N) * E);
        char *dst = reinterpret_cast<char *>(cbuf + (size_t)buf * N * E);
        for (size_t off = 0; off < c_bytes; off += 16384)
            tma_load_1d(dst + off, src + off, (unsigned)(c_bytes - off < 16384 ? c_bytes - off : 16384), &bar[buf]);
    };
    if (HAS_C && tid == 0 && (int64_t)blockIdx.x < a.B) stage_context(blockIdx.x, 0);

    int it = 0;
    for (int64_t b = blockIdx.x; b < a.B; b += gridDim.x, ++it) {
        const int64_t r = b;                        // single start: rollout index = instance
        const int buf = it & 1;
        const float *C = cbuf + (size_t)buf * N * E;
        // ------------------------------------------------ registers <- tables (global / L2)
        u64 K2[M][4], V2[M][4], L2[M][4];
        {
            const float *kb = a.tables + ((size_t)0 * Mrows + b * (int64_t)N) * E + dim0;
            const float *vb = a.tables + ((size_t)1 * Mrows + b * (int64_t)N) * E + dim0;
            const float *lb = a.tables + ((size_t)2 * Mrows + b * (int64_t)N) * E + dim0;
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const int n = g + 16 * j;
                if (n < N) {
                    const ulonglong2 k0 = __ldg(reinterpret_cast<const ulonglong2 *>(kb + (size_t)n * E));
                    const ulonglong2 k1 = __ldg(reinterpret_cast<const ulonglong2 *>(kb + (size_t)n * E) + 1);
                    const ulonglong2 v0 = __ldg(reinterpret_cast<const ulonglong2 *>(vb + (size_t)n * E));
                    const ulonglong2 v1 = __ldg(reinterpret_cast<const ulonglong2 *>(vb + (size_t)n * E) + 1);
                    const ulonglong2 l0 = __ldg(reinterpret_cast<const ulonglong2 *>(lb + (size_t)n * E));
                    const ulonglong2 l1 = __ldg(reinterpret_cast<const ulonglong2 *>(lb + (size_t)n * E) + 1);
                    K2[j][0] = k0.x; K2[j][1] = k0.y; K2[j][2] = k1.x; K2[j][3] = k1.y;
                    V2[j][0] = v0.x; V2[j][1] = v0.y; V2[j][2] = v1.x; V2[j][3] = v1.y;
                    L2[j][0] = l0.x; L2[j][1] = l0.y; L2[j][2] = l1.x; L2[j][3] = l1.y;
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) K2[j][i] = V2[j][i] = L2[j][i] = 0ull;
                }
            }
        }
        // constant part of the query, pre-scaled: qbs = scale * (gctx [+ step-0 placeholder])
        u64 qbs2[4], wcs2[4];
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
        set_qb(ENV == RL4CO_ENV_TSP ? a.q0 + dim0 : nullptr);   // tsp step 0: context.py:111-121
        if (ENV == RL4CO_ENV_CVRP) {                              // remaining-capacity column, context.py:317-331
            const float4 w0 = __ldg(reinterpret_cast<const float4 *>(a.wcap + dim0));
            const float4 w1 = __ldg(reinterpret_cast<const float4 *>(a.wcap + dim0) + 1);
            wcs2[0] = pack2(w0.x * kScoreScale, w0.y * kScoreScale);
            wcs2[1] = pack2(w0.z * kScoreScale, w0.w * kScoreScale);
            wcs2[2] = pack2(w1.x * kScoreScale, w1.y * kScoreScale);
            wcs2[3] = pack2(w1.z * kScoreScale, w1.w * kScoreScale);
        }
        // ------------------------------------------------ small per-instance state -> smem
        __syncthreads();                            // previous instance's epilogue is done with smem
        if (tid < NS) {
            sxs[tid] = -INFINITY;
            s_dem[tid] = (ENV == RL4CO_ENV_CVRP && tid >= 1 && tid < N) ? a.demand[b * (int64_t)(N - 1) + tid - 1] : 0.f;
        }
        if (tid < kWarps) scand[tid] = make_float2(-INFINITY, 0.f);
        {
            const float2 *lrow = reinterpret_cast<const float2 *>(a.locs + b * (int64_t)N * 2);
            for (int n = tid; n < N; n += kThreads) s_locs[n] = lrow[n];
            if (a.mode == RL4CO_MODE_EVALUATE)
                for (int t = tid; t < T; t += kThreads) {
                    int64_t gv = a.actions_in[(size_t)r * T + t];
                    s_ain[t] = (gv < 0 || gv >= N) ? -1 : (int)gv;
                }
        }
        // next instance: context table -> the other buffer (its previous user finished before
        // the barrier above), K/V/L rows -> L2
        if (tid == 0 && b + gridDim.x < a.B) {
            const int64_t nb = b + gridDim.x;
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
        // bit j of a word <-> node g + 16 j.  tsp/env.py:95-120, cvrp/env.py:107-135, slap/env.py:95-129
        unsigned own = 0;
#pragma unroll
        for (int j = 0; j < M; ++j) own |= (g + 16 * j < N) ? (1u << j) : 0u;
        unsigned alive = own, vis = 0;
        if (ENV == RL4CO_ENV_SLAP && g == 0) alive &= ~1u;      // slap: depot never assignable
        float dm[M];
#pragma unroll
        for (int j = 0; j < M; ++j) dm[j] = (ENV == RL4CO_ENV_CVRP) ? s_dem[g + 16 * j] : 0.f;
        const float cap = (ENV == RL4CO_ENV_CVRP) ? a.vehicle_capacity[b] : 0.f;
        float used = 0.f;
        auto cvrp_mask = [&](int at) {              // cvrp/env.py:137-149
            unsigned f = 0;
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const bool exceeds = __fadd_rn(dm[j], used) > cap;
                if (!exceeds && !((vis >> j) & 1u)) f |= 1u << j;
            }
            f &= own;
            if (g == 0) f &= ~1u;
            const bool any = __any_sync(kFull, f != 0);
            if (g == 0 && !(at == 0 && any)) f |= 1u;
            return f;
        };
        if (ENV == RL4CO_ENV_CVRP) alive = cvrp_mask(0);
        int cur = 0, t = 0, status = 0;
        bool done = false, use_c = (ENV == RL4CO_ENV_CVRP);
        const int t_end = (ENV == RL4CO_ENV_TSP) ? (N < T ? N : T) : (ENV == RL4CO_ENV_SLAP ? (a.P < T ? a.P : T) : T);

        if (HAS_C) {                                // context table landed?
            mbar_wait(&bar[buf], parity[buf]);
            parity[buf] ^= 1;
        }

        // ================================================== decode loop ==================
        while (!done && t < t_end) {
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
            if (ENV == RL4CO_ENV_CVRP) {
                const float rem = __fsub_rn(cap, used);
                const u64 rem2 = pack2(rem, rem);
#pragma unroll
                for (int i = 0; i < 4; ++i) q2[i] = ffma2(wcs2[i], rem2, q2[i]);
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
#pragma unroll
                for (int j = 0; j < M; ++j) s[j] = hsum2(acc[j]);
#pragma unroll
                for (int j = 0; j < M; ++j) s[j] += __shfl_xor_sync(kFull, s[j], 16);
            }
            float mx = -INFINITY;
#pragma unroll
            for (int j = 0; j < M; ++j) {
                s[j] = ((alive >> j) & 1u) ? s[j] : -INFINITY;
                mx = fmaxf(mx, s[j]);
            }
            mx = redux_max(mx);
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
                    w4[k] = keep + __shfl_xor_sync(kFull, send, 8);
                }
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const float send = b2 ? w4[k] : w4[k + 2], keep = b2 ? w4[k + 2] : w4[k];
                    w2[k] = keep + __shfl_xor_sync(kFull, send, 4);
                }
                {
                    const float send = b1 ? w2[0] : w2[1], keep = b1 ? w2[1] : w2[0];
                    w1 = keep + __shfl_xor_sync(kFull, send, 2);
                }
                w1 += __shfl_xor_sync(kFull, w1, 1);
                // this lane now holds head dim (b3 ? 4 : 0) + (b2 ? 2 : 0) + (b1 ? 1 : 0) of its half
                const int d = (b3 ? 4 : 0) + (b2 ? 2 : 0) + (b1 ? 1 : 0);
                so[dim0 + d] = __fdividef(w1, psum);
            }
            __syncwarp();
            // ---- partial pointer logits over this half head, project_out folded (attention.py:276-282)
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
                float *row = spl + (2 * warp + hh) * PLS + g;
#pragma unroll
                for (int j = 0; j < M; ++j) row[16 * j] = hsum2(acc[j]);
            }
            __syncthreads();

            // ---- warp w < M finishes nodes 16w .. 16w+15: process_logits (utils/decoding.py:141-191)
            float x = -INFINITY;
            if (warp < M) {
                const float *col = spl + hh * PLS + 16 * warp + g;
                float sum = 0.f;
#pragma unroll
                for (int k = 0; k < 8; ++k) sum += col[2 * k * PLS];
                sum += __shfl_xor_sync(kFull, sum, 16);
                const bool exists = (own >> warp) & 1u, feasible = (alive >> warp) & 1u;
                if (exists && (feasible || !a.mask_logits)) {
                    x = sum * kInvSqrtE;
                    if (x != x) status |= RL4CO_ST_NAN_LOGITS;
                    if (a.tanh_clipping > 0.f) x = tanhf(x) * a.tanh_clipping;
                    if (a.temperature != 1.f) x = __fdiv_rn(x, a.temperature);
                }
                const float mw = redux_max(x);
                if (a.mode != RL4CO_MODE_GREEDY && hh == 0) sxs[16 * warp + g] = x;
                float e = (x == -INFINITY) ? 0.f : expf(x - mw);
                if (!sampling) {
                    const unsigned eq = __ballot_sync(kFull, x == mw) & 0xffffu;
                    const int arg = 16 * warp + (eq ? __ffs(eq) - 1 : 0);
                    if (lane == 0) scand[warp] = make_float2(mw, __int_as_float(arg));
#pragma unroll
                    for (int o = 8; o > 0; o >>= 1) e += __shfl_xor_sync(kFull, e, o);
                } else {
                    // inclusive prefix over the 16 nodes (both halves hold the same values)
#pragma unroll
                    for (int o = 1; o < 16; o <<= 1) {
                        const float up = __shfl_up_sync(kFull, e, o, 16);
                        if (g >= o) e += up;
                    }
                    if (hh == 0) spref[16 * warp + g] = e;
                    e = __shfl_sync(kFull, e, 15);
                    if (lane == 0) scand[warp] = make_float2(mw, e);
                }
                if (lane == 0) sring[(size_t)t * kWarps + warp] = make_float2(mw, e);
            } else if (lane == 0) {
                sring[(size_t)t * kWarps + warp] = make_float2(-INFINITY, 0.f);
            }
            __syncthreads();

            // ---- every warp: combine the candidates, select (utils/decoding.py:371-397,410-499)
            const float2 cd = scand[lane & 7];
            const float gm = redux_max(cd.x);
            int act;
            float xa;
            if (a.mode == RL4CO_MODE_GREEDY) {
                // argmax, lowest index on ties (the log-softmax shift is common to all nodes)
                const unsigned eq = __ballot_sync(kFull, cd.x == gm) & 0xffu;
                act = __float_as_int(__shfl_sync(kFull, cd.y, eq ? __ffs(eq) - 1 : 0));
                xa = gm;
            } else {
                if (a.mode == RL4CO_MODE_EVALUATE) {
                    act = s_ain[t];
                    if (act < 0) { status |= RL4CO_ST_ACTION_RANGE; act = 0; }
                } else {
                    // inverse-CDF sample of Categorical(exp(logp)) with one Philox uniform
                    const float sc = (cd.x == -INFINITY) ? 0.f : expf(cd.x - gm);
                    const float tw = cd.y * sc;
                    float cum = tw;
#pragma unroll
                    for (int o = 1; o < 8; o <<= 1) {
                        const float up = __shfl_up_sync(kFull, cum, o, 8);
                        if ((lane & 7) >= o) cum += up;
                    }
                    const float total = __shfl_sync(kFull, cum, 7);
                    const float u = philox_uniform(a.seed, a.offset, (uint64_t)r, (uint32_t)t) * total;
                    const unsigned over = __ballot_sync(kFull, cum > u && tw > 0.f) & 0xffu;
                    const unsigned nz = __ballot_sync(kFull, tw > 0.f) & 0xffu;
                    const int ws = over ? __ffs(over) - 1 : (nz ? 31 - __clz(nz) : 0);
                    const float before = __shfl_sync(kFull, cum - tw, ws), scw = __shfl_sync(kFull, sc, ws);
                    const float pf = spref[16 * ws + g];
                    const float prev = __shfl_up_sync(kFull, pf, 1, 16);
                    const bool nonzero = pf > (g == 0 ? 0.f : prev);
                    const unsigned hit = __ballot_sync(kFull, nonzero && fmaf(pf, scw, before) > u) & 0xffffu;
                    const unsigned any = __ballot_sync(kFull, nonzero) & 0xffffu;
                    act = 16 * ws + (hit ? __ffs(hit) - 1 : (any ? 31 - __clz(any) : 0));
                }
                xa = sxs[act];
            }
            // feasibility of the chosen action (utils/decoding.py:377-379,393-395)
            {
                const unsigned fb = __shfl_sync(kFull, alive, act & 15);
                if (!((fb >> (act >> 4)) & 1u)) status |= RL4CO_ST_INFEASIBLE_ACTION;
            }
            if (tid == 0) { s_act[t] = act; s_xa[t] = xa; }
            // ---- transition
            const unsigned abit = (g == (act & 15)) ? (1u << (act >> 4)) : 0u;
            if (ENV == RL4CO_ENV_TSP) {                 // tsp/env.py:67-93
                alive &= ~abit;
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
                alive = cvrp_mask(act);
                done = __all_sync(kFull, vis == own);
            } else {                                    // slap/env.py:38-93
                alive &= ~abit;
            }
            cur = act;
            ++t;
        }

        // ================================================== epilogue =====================
        __syncthreads();
        const int nact = t;
        // chosen log-probs from the per-step ingredients: logp = (x_a - m) - log(sum exp(x - m))
        for (int q = tid; q < nact; q += kThreads) {
            const float2 *rg = sring + (size_t)q * kWarps;
            float m = -INFINITY;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) m = fmaxf(m, rg[w].x);
            float tot = 0.f;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                const float2 c = rg[w];
                tot += (c.x == -INFINITY) ? 0.f : c.y * expf(c.x - m);
            }
            const float lpa = (s_xa[q] - m) - logf(tot);
            s_lp[q] = lpa;
            if (!(lpa > -1000.f)) status |= RL4CO_ST_NEG_INF_LOGP;
        }
        float part = 0.f;
        if (ENV == RL4CO_ENV_TSP) {              // tsp/env.py:157-163
            for (int q = tid; q < nact; q += kThreads) {
                float2 p = s_locs[s_act[q]], w = s_locs[s_act[q + 1 == nact ? 0 : q + 1]];
                part += leg(p.x, p.y, w.x, w.y);
            }
        } else if (ENV == RL4CO_ENV_CVRP) {      // cvrp/env.py:151-160 (depot at both ends)
            for (int q = tid; q <= nact; q += kThreads) {
                float2 p = q == 0 ? s_locs[0] : s_locs[s_act[q - 1]];
                float2 w = q == nact ? s_locs[0] : s_locs[s_act[q]];
                part += leg(p.x, p.y, w.x, w.y);
            }
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
    const Layout L = make_layout(a.N, a.T, M, ENV != RL4CO_ENV_SLAP);
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
    const int64_t grid = a.B < sm_count ? a.B : sm_count;   // persistent: one CTA per SM
    am_rollout_rf_kernel<ENV, M><<<(unsigned)grid, kThreads, L.total, stream>>>(a);
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
    if (a.S > 1 || a.multistart || a.N > 128 || a.logp_all != nullptr) return false;
    return rf::make_layout(a.N, a.T, rf::nodes_per_thread(a.N), a.env != RL4CO_ENV_SLAP).total <= (size_t)227 * 1024;
}

int launch_rollout_rf(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    switch (a.env) {
        case RL4CO_ENV_TSP: return rf::launch_env<RL4CO_ENV_TSP>(a, stream);
        case RL4CO_ENV_CVRP: return rf::launch_env<RL4CO_ENV_CVRP>(a, stream);
        default: return rf::launch_env<RL4CO_ENV_SLAP>(a, stream);
    }
}

}  // namespace rl4co
