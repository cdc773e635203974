// K2 (single start) - persistent Attention-Model rollout with REGISTER-FILE resident tables.
//
// One CTA of 512 threads per SM, one instance at a time, the whole trajectory inside the
// kernel.  The instance's three per-node tables (glimpse K, glimpse V, folded logit K:
// 3 x N x 128 fp32 = 150 KB at N=100) live in the register file for the whole trajectory,
// distributed over the 512 threads so that every per-step mat-vec is thread-local FMA work
// on registers (shared-memory bandwidth would cap the step at 3*N*E*4 B / 128 B/clk):
//
//   role A  (node-major, K and L):  warp w owns the node group [w*G, w*G+G); lane = (c, i)
//           with c = lane&3 the 32-dim chunk (heads 2c, 2c+1) and i = lane>>2 the node of
//           the group.  32 K + 32 L registers per thread.
//   role B  (dim-major, V):  warp w owns dims [8w, 8w+8) (half of head w>>1); lane = (j, dd)
//           with j = lane&3 the node quarter [4G*j, 4G*j+4G) and dd = lane>>2 the dim.
//           4G V registers per thread.
//   role C  (selection, every warp redundantly): lane l owns nodes 4l..4l+3 of the logit
//           vector and of the environment state (feasible / visited bits), so the action,
//           the transition and the new mask are known to all warps without a barrier.
//
// One decode step = 3 block barriers:
//   A: scores s[h][n] = sb[h][n] + <C[cur]_h, K[n]_h>   (C = per-node context table, the one
//      table that is gathered by row, staged in shared memory by a TMA bulk copy),
//      group-local max, p = 2^(s' - m_group) -> smem                          | barrier 1
//   B: o[d] = sum_g 2^(m_g - M) sum_{n in g} p[n] V[n][d] / (same for 1), quarters combined
//      with two shuffles -> smem                                               | barrier 2
//   A: logit[n] = <o, L[n]> (project_out is folded into L), /sqrt(E), tanh clip, mask,
//      temperature -> smem                                                     | barrier 3
//   C: argmax / Philox inverse-CDF sample / given action, log-softmax for the chosen
//      log-prob (one warp per step, round robin), env transition + new mask (all warps).
//
// Reference op order being reproduced (all fp32): SURVEY.md appendix A;
// rl4co/models/zoo/am/decoder.py:134-200, rl4co/models/nn/attention.py:263-309,
// rl4co/utils/decoding.py:141-191,327-397, rl4co/envs/routing/{tsp,cvrp}/env.py,
// rl4co/envs/warehousing/slap/env.py:38-143.
#include "rollout_util.cuh"

namespace rl4co {
namespace rf {

constexpr int E = RL4CO_EMBED_DIM;   // 128
constexpr int kThreads = 512;
constexpr int kWarps = 16;
constexpr int kSlots = 128;           // node slots of the selection layout (4 per lane)
constexpr float kScoreScale = 0.25f * 1.4426950408889634f;   // 1/sqrt(16) * log2(e)
constexpr float kTieWindow = 2e-6f;   // greedy: candidates this close to the max go through
                                      // the exact log-softmax comparison (see select())

__host__ __device__ constexpr int quarter_stride(int G) {      // floats between node quarters
    return ((4 * G) % 32 == 0 || (4 * G) % 32 == 16) ? 4 * G + 4 : 4 * G;
}
__host__ __device__ constexpr int head_stride(int G) { return 4 * quarter_stride(G) + 4; }

struct Layout {
    size_t off_c, off_p, off_mw, off_o, off_qb, off_x, off_act, off_lp, off_ain, off_locs,
        off_dem, off_part, off_bar, total;
};
__host__ __device__ inline size_t up16(size_t v) { return (v + 15) & ~(size_t)15; }
__host__ __device__ inline Layout make_layout(int N, int T, int G, bool has_c) {
    Layout L;
    size_t o = 0;
    L.off_c = o;    o += has_c ? (size_t)2 * N * E * 4 : 0;     // double-buffered context table
    L.off_p = o;    o += (size_t)8 * head_stride(G) * 4;
    L.off_mw = o;   o += 8 * 16 * 4;
    L.off_o = o;    o += E * 4;
    L.off_qb = o;   o += E * 4;
    L.off_x = o;    o += kSlots * 4;
    L.off_act = o;  o += up16((size_t)T * 4);
    L.off_lp = o;   o += up16((size_t)T * 4);
    L.off_ain = o;  o += up16((size_t)T * 4);
    L.off_locs = o; o += up16((size_t)N * 8);
    L.off_dem = o;  o += kSlots * 4;
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
// max over the lanes named by `mask` (all of them execute this with the same mask)
__device__ __forceinline__ float group_max(float v, unsigned mask) {
    float r;
    asm volatile("redux.sync.max.f32 %0, %1, %2;" : "=f"(r) : "f"(v), "r"(mask));
    return r;
}
__device__ __forceinline__ float dot4(float4 a, float4 b, float s) {
    s = fmaf(a.x, b.x, s); s = fmaf(a.y, b.y, s); s = fmaf(a.z, b.z, s); s = fmaf(a.w, b.w, s);
    return s;
}
// float4 slot k of a 32-float chunk, rotated by the chunk index inside each head half so that
// the four chunks read by one warp instruction hit four different bank groups
__device__ __forceinline__ int rot(int k, int c) { return (k & 4) | ((k + c) & 3); }

template <int ENV, int G>
__global__ void __launch_bounds__(kThreads, 1)
am_rollout_rf_kernel(const rl4co_am_rollout_args a) {
    constexpr int NB = 4 * G;                       // nodes per role-B thread
    constexpr int QS = quarter_stride(G), PS = head_stride(G);
    constexpr bool HAS_C = ENV != RL4CO_ENV_SLAP;
    extern __shared__ __align__(128) unsigned char smem[];
    const int N = a.N, T = a.T;
    const Layout L = make_layout(N, T, G, HAS_C);
    float *cbuf = reinterpret_cast<float *>(smem + L.off_c);
    float *sp = reinterpret_cast<float *>(smem + L.off_p);
    float *smw = reinterpret_cast<float *>(smem + L.off_mw);
    float *so = reinterpret_cast<float *>(smem + L.off_o);
    float *sqb = reinterpret_cast<float *>(smem + L.off_qb);
    float *sx = reinterpret_cast<float *>(smem + L.off_x);
    int *s_act = reinterpret_cast<int *>(smem + L.off_act);
    float *s_lp = reinterpret_cast<float *>(smem + L.off_lp);
    int *s_ain = reinterpret_cast<int *>(smem + L.off_ain);
    float2 *s_locs = reinterpret_cast<float2 *>(smem + L.off_locs);
    float *s_dem = reinterpret_cast<float *>(smem + L.off_dem);     // demand of node n (0 for depot)
    float *spart = reinterpret_cast<float *>(smem + L.off_part);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + L.off_bar); // one per context buffer

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t M = a.B * (int64_t)N;             // rows per table
    // role A
    const int ac = lane & 3, ai = lane >> 2;
    const int an = warp * G + ai;                   // own node
    const bool a_slot = ai < G;                     // lane takes part in role A
    const bool a_valid = a_slot && an < N;
    const unsigned a_grp = 0x11111111u << ac;       // lanes with the same chunk
    // role B
    const int bj = lane & 3, bd = warp * 8 + (lane >> 2), bh = warp >> 1;
    const float sqrt_e = sqrtf((float)E);
    const size_t c_bytes = (size_t)N * E * 4;

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_barrier_init();
    }
    __syncthreads();
    unsigned parity[2] = {0u, 0u};
    // context table of the first instance of this CTA
    if (HAS_C && tid == 0 && (int64_t)blockIdx.x < a.B) {
        mbar_expect_tx(&bar[0], (unsigned)c_bytes);
        const char *src = reinterpret_cast<const char *>(a.tables + ((size_t)3 * M + (int64_t)blockIdx.x * N) * E);
        for (size_t off = 0; off < c_bytes; off += 16384)
            tma_load_1d(reinterpret_cast<char *>(cbuf) + off, src + off,
                        (unsigned)(c_bytes - off < 16384 ? c_bytes - off : 16384), &bar[0]);
    }

    int it = 0;
    for (int64_t b = blockIdx.x; b < a.B; b += gridDim.x, ++it) {
        const int64_t r = b;                        // single start: rollout index = instance
        const int buf = it & 1;
        float *C = cbuf + (size_t)buf * N * E;
        // ------------------------------------------------ registers <- tables (global)
        float4 kreg[8], lreg[8];
        float vreg[NB];
        {
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 *kp = reinterpret_cast<const float4 *>(a.tables + ((size_t)0 * M + b * (int64_t)N + an) * E + 32 * ac);
            const float4 *lp = reinterpret_cast<const float4 *>(a.tables + ((size_t)2 * M + b * (int64_t)N + an) * E + 32 * ac);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                kreg[k] = a_valid ? __ldg(kp + rot(k, ac)) : z;
                lreg[k] = a_valid ? __ldg(lp + rot(k, ac)) : z;
            }
            const float *vp = a.tables + ((size_t)1 * M + b * (int64_t)N) * E + bd;
#pragma unroll
            for (int k = 0; k < NB; ++k) {
                const int n = NB * bj + k;
                vreg[k] = n < N ? __ldg(vp + (size_t)n * E) : 0.f;
            }
        }
        // ------------------------------------------------ small per-instance state -> smem
        __syncthreads();                            // previous instance's epilogue is done with smem
        if (tid < E) {
            float g = a.gctx ? a.gctx[b * E + tid] : 0.f;
            if (ENV == RL4CO_ENV_TSP) g += a.q0[tid];           // step-0 context (context.py:111-121)
            sqb[tid] = g;
        }
        if (tid < kSlots) {
            sx[tid] = -INFINITY;
            if (ENV == RL4CO_ENV_CVRP)
                s_dem[tid] = (tid >= 1 && tid < N) ? a.demand[b * (int64_t)(N - 1) + tid - 1] : 0.f;
        }
        {
            const float2 *lrow = reinterpret_cast<const float2 *>(a.locs + b * (int64_t)N * 2);
            for (int n = tid; n < N; n += kThreads) s_locs[n] = lrow[n];
            if (a.mode == RL4CO_MODE_EVALUATE)
                for (int t = tid; t < T; t += kThreads) {
                    int64_t g = a.actions_in[(size_t)r * T + t];
                    s_ain[t] = (g < 0 || g >= N) ? -1 : (int)g;
                }
        }
        // next instance's context table -> the other buffer (its previous user finished
        // before the barrier above)
        if (HAS_C && tid == 0 && b + gridDim.x < a.B) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&bar[buf ^ 1], (unsigned)c_bytes);
            const char *src = reinterpret_cast<const char *>(a.tables + ((size_t)3 * M + (b + gridDim.x) * (int64_t)N) * E);
            char *dst = reinterpret_cast<char *>(cbuf + (size_t)(buf ^ 1) * N * E);
            for (size_t off = 0; off < c_bytes; off += 16384)
                tma_load_1d(dst + off, src + off,
                            (unsigned)(c_bytes - off < 16384 ? c_bytes - off : 16384), &bar[buf ^ 1]);
        }
        __syncthreads();

        // constant part of the scores: sb[h][n] = <qb_h, K[n]_h>
        float sb0, sb1, kw0 = 0.f, kw1 = 0.f;
        auto dot_k = [&](const float *vec, float &r0, float &r1) {
            const float4 *q4 = reinterpret_cast<const float4 *>(vec + 32 * ac);
            float x0 = 0.f, x1 = 0.f, y0 = 0.f, y1 = 0.f;
#pragma unroll
            for (int k = 0; k < 4; k += 2) {
                x0 = dot4(q4[rot(k, ac)], kreg[k], x0);
                x1 = dot4(q4[rot(k + 1, ac)], kreg[k + 1], x1);
                y0 = dot4(q4[rot(k + 4, ac)], kreg[k + 4], y0);
                y1 = dot4(q4[rot(k + 5, ac)], kreg[k + 5], y1);
            }
            r0 = x0 + x1;
            r1 = y0 + y1;
        };
        dot_k(sqb, sb0, sb1);
        if (ENV == RL4CO_ENV_CVRP) {                // remaining-capacity column (context.py:317-331)
            __syncthreads();
            if (tid < E) sqb[tid] = a.wcap[tid];
            __syncthreads();
            dot_k(sqb, kw0, kw1);
        }

        // ------------------------------------------------ env state (role C layout)
        // tsp/env.py:95-120, cvrp/env.py:107-135, slap/env.py:95-129
        unsigned own = 0;                           // bits of the nodes 4l..4l+3 that exist
#pragma unroll
        for (int k = 0; k < 4; ++k) own |= (4 * lane + k < N) ? (1u << k) : 0u;
        unsigned feas = own, vis = 0;
        if (ENV == RL4CO_ENV_SLAP && lane == 0) feas &= ~1u;    // slap: depot never assignable
        const float cap = (ENV == RL4CO_ENV_CVRP) ? a.vehicle_capacity[b] : 0.f;
        float used = 0.f;
        // cvrp/env.py:137-149 on the role-C layout (every warp computes the same bits)
        auto cvrp_mask = [&](int at) {
            const float4 d4 = *reinterpret_cast<const float4 *>(s_dem + 4 * lane);
            const float dv[4] = {d4.x, d4.y, d4.z, d4.w};
            unsigned f = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const bool exceeds = __fadd_rn(dv[k], used) > cap;
                if (!exceeds && !((vis >> k) & 1u)) f |= 1u << k;
            }
            f &= own;
            if (lane == 0) f &= ~1u;
            const bool any = __any_sync(kFull, f != 0);
            if (lane == 0 && !(at == 0 && any)) f |= 1u;
            return f;
        };
        if (ENV == RL4CO_ENV_CVRP) feas = cvrp_mask(0);
        int cur = 0, t = 0, status = 0;
        bool done = false, use_c = (ENV == RL4CO_ENV_CVRP);
        const int t_end = (ENV == RL4CO_ENV_TSP) ? (N < T ? N : T) : (ENV == RL4CO_ENV_SLAP ? (a.P < T ? a.P : T) : T);

        if (HAS_C) {                                // context table landed?
            mbar_wait(&bar[buf], parity[buf]);
            parity[buf] ^= 1;
        }

        // ================================================== decode loop ==================
        while (!done && t < t_end) {
            // ---- role A: glimpse scores (attention.py:289-306), group max, exponentials
            const unsigned own_bits = __shfl_sync(kFull, feas, an >> 2);    // every lane takes part
            const bool alive = a_valid && ((own_bits >> (an & 3)) & 1u);
            {
                float x0 = sb0, x1 = 0.f, y0 = sb1, y1 = 0.f;
                if (HAS_C && use_c) {
                    const float4 *c4 = reinterpret_cast<const float4 *>(C + (size_t)cur * E + 32 * ac);
#pragma unroll
                    for (int k = 0; k < 4; k += 2) {
                        x0 = dot4(c4[rot(k, ac)], kreg[k], x0);
                        x1 = dot4(c4[rot(k + 1, ac)], kreg[k + 1], x1);
                        y0 = dot4(c4[rot(k + 4, ac)], kreg[k + 4], y0);
                        y1 = dot4(c4[rot(k + 5, ac)], kreg[k + 5], y1);
                    }
                }
                float s0 = x0 + x1, s1 = y0 + y1;
                if (ENV == RL4CO_ENV_CVRP) {
                    const float rem = __fsub_rn(cap, used);
                    s0 = fmaf(rem, kw0, s0);
                    s1 = fmaf(rem, kw1, s1);
                }
                s0 = alive ? s0 * kScoreScale : -INFINITY;
                s1 = alive ? s1 * kScoreScale : -INFINITY;
                const float m0 = group_max(s0, a_grp), m1 = group_max(s1, a_grp);
                const float p0 = alive ? ex2f(s0 - m0) : 0.f;
                const float p1 = alive ? ex2f(s1 - m1) : 0.f;
                if (a_slot) {
                    const int pos = an + (QS - NB) * (an / NB);
                    sp[(2 * ac) * PS + pos] = p0;
                    sp[(2 * ac + 1) * PS + pos] = p1;
                    if (ai == 0) {
                        smw[(2 * ac) * 16 + warp] = m0;
                        smw[(2 * ac + 1) * 16 + warp] = m1;
                    }
                }
            }
            __syncthreads();

            // ---- role B: heads o = softmax . V with per-group rescaling
            {
                const float4 mg = *reinterpret_cast<const float4 *>(smw + bh * 16 + 4 * bj);
                float mx = fmaxf(fmaxf(mg.x, mg.y), fmaxf(mg.z, mg.w));
                mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, 1));
                mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, 2));
                float sc[4];
                sc[0] = mg.x == -INFINITY ? 0.f : ex2f(mg.x - mx);
                sc[1] = mg.y == -INFINITY ? 0.f : ex2f(mg.y - mx);
                sc[2] = mg.z == -INFINITY ? 0.f : ex2f(mg.z - mx);
                sc[3] = mg.w == -INFINITY ? 0.f : ex2f(mg.w - mx);
                const float4 *p4 = reinterpret_cast<const float4 *>(sp + bh * PS + bj * QS);
                float acc[4] = {0.f, 0.f, 0.f, 0.f}, ls[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int q = 0; q < G; ++q) {
                    const float4 pv = p4[q];
                    const float pe[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int k = 4 * q + e;             // node k of this thread, group k / G
                        acc[k / G] = fmaf(pe[e], vreg[k], acc[k / G]);
                        ls[k / G] += pe[e];
                    }
                }
                float osum = sc[0] * acc[0], lsum = sc[0] * ls[0];
#pragma unroll
                for (int g = 1; g < 4; ++g) {
                    osum = fmaf(sc[g], acc[g], osum);
                    lsum = fmaf(sc[g], ls[g], lsum);
                }
                osum += __shfl_xor_sync(kFull, osum, 1);
                lsum += __shfl_xor_sync(kFull, lsum, 1);
                osum += __shfl_xor_sync(kFull, osum, 2);
                lsum += __shfl_xor_sync(kFull, lsum, 2);
                if (bj == 0) so[bd] = __fdividef(osum, lsum);
            }
            __syncthreads();

            // ---- role A: pointer logits with folded project_out (attention.py:276-285), then
            //      process_logits (utils/decoding.py:141-191): tanh clip, mask, temperature
            {
                const float4 *o4 = reinterpret_cast<const float4 *>(so + 32 * ac);
                float x0 = 0.f, x1 = 0.f, y0 = 0.f, y1 = 0.f;
#pragma unroll
                for (int k = 0; k < 4; k += 2) {
                    x0 = dot4(o4[rot(k, ac)], lreg[k], x0);
                    x1 = dot4(o4[rot(k + 1, ac)], lreg[k + 1], x1);
                    y0 = dot4(o4[rot(k + 4, ac)], lreg[k + 4], y0);
                    y1 = dot4(o4[rot(k + 5, ac)], lreg[k + 5], y1);
                }
                float p = (x0 + x1) + (y0 + y1);
                p += __shfl_xor_sync(kFull, p, 1);
                p += __shfl_xor_sync(kFull, p, 2);
                if (ac == 0 && a_valid) {
                    float x = -INFINITY;
                    if (alive || !a.mask_logits) {
                        x = __fdiv_rn(p, sqrt_e);
                        if (isnan(x)) status |= RL4CO_ST_NAN_LOGITS;
                        if (a.tanh_clipping > 0.f) x = tanhf(x) * a.tanh_clipping;
                        x = __fdiv_rn(x, a.temperature);
                    }
                    sx[an] = x;
                }
            }
            __syncthreads();

            // ---- role C (every warp, same registers): selection + transition
            const float4 x4 = *reinterpret_cast<const float4 *>(sx + 4 * lane);
            const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
            const float m = group_max(fmaxf(fmaxf(xv[0], xv[1]), fmaxf(xv[2], xv[3])), kFull);
            const bool lp_warp = (warp == (t & (kWarps - 1)));      // writes this step's log-prob
            int act = -1;
            bool need_lse = lp_warp || a.logp_all != nullptr;
            if (a.mode == RL4CO_MODE_GREEDY) {
                // argmax of the log-probs, lowest index on ties (utils/decoding.py:371-381).
                // log_softmax subtracts the same constant from every logit, so the argmax of
                // the logits is the answer unless rounding of that subtraction merges values:
                // only then (>1 candidate within kTieWindow) compare the log-probs themselves.
                const float thr = m - kTieWindow;
                const unsigned b0 = __ballot_sync(kFull, xv[0] >= thr), b1 = __ballot_sync(kFull, xv[1] >= thr),
                               b2 = __ballot_sync(kFull, xv[2] >= thr), b3 = __ballot_sync(kFull, xv[3] >= thr);
                if (__popc(b0) + __popc(b1) + __popc(b2) + __popc(b3) == 1)
                    act = b0 ? 4 * (__ffs(b0) - 1) : b1 ? 4 * (__ffs(b1) - 1) + 1
                        : b2 ? 4 * (__ffs(b2) - 1) + 2 : 4 * (__ffs(b3) - 1) + 3;
                else
                    need_lse = true;
            } else if (a.mode == RL4CO_MODE_EVALUATE) {
                act = s_ain[t];
                if (act < 0) { status |= RL4CO_ST_ACTION_RANGE; act = 0; }
            } else {
                need_lse = true;
            }
            if (need_lse) {
                float e[4], lsum = 0.f;
#pragma unroll
                for (int k = 0; k < 4; ++k) { e[k] = (xv[k] == -INFINITY) ? 0.f : expf(xv[k] - m); lsum += e[k]; }
                const float total = warp_sum(lsum);
                const float lse = logf(total);
                if (a.mode == RL4CO_MODE_GREEDY && act < 0) {
                    float bv = (xv[0] - m) - lse; int bi = lane * 4;
#pragma unroll
                    for (int k = 1; k < 4; ++k) { float v = (xv[k] - m) - lse; if (v > bv) { bv = v; bi = lane * 4 + k; } }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        float ov = __shfl_xor_sync(kFull, bv, o);
                        int oi = __shfl_xor_sync(kFull, bi, o);
                        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
                    }
                    act = bi;
                } else if (a.mode == RL4CO_MODE_SAMPLING) {
                    // inverse-CDF sample of Categorical(exp(logp)) with a Philox uniform
                    float u = philox_uniform(a.seed, a.offset, (uint64_t)r, (uint32_t)t) * total;
                    float incl = lsum;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        float up = __shfl_up_sync(kFull, incl, o);
                        if (lane >= o) incl += up;
                    }
                    float run = incl - lsum;
                    int pick = -1, lastf = -1;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        if (e[k] > 0.f) {
                            lastf = lane * 4 + k;
                            run += e[k];
                            if (pick < 0 && run > u) pick = lane * 4 + k;
                        }
                    }
                    unsigned has = __ballot_sync(kFull, pick >= 0);
                    int lastall = lastf;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) lastall = max(lastall, __shfl_xor_sync(kFull, lastall, o));
                    act = has ? __shfl_sync(kFull, pick, __ffs(has) - 1) : lastall;
                }
                if (lp_warp || a.logp_all) {
                    const int ak = act & 3;
                    const float mine = ak == 0 ? xv[0] : ak == 1 ? xv[1] : ak == 2 ? xv[2] : xv[3];
                    const float xa = __shfl_sync(kFull, mine, act >> 2);
                    const float lpa = (xa - m) - lse;
                    if (lp_warp) {
                        if (lane == 0) {
                            s_lp[t] = lpa;
                            if (!(lpa > -1000.f)) status |= RL4CO_ST_NEG_INF_LOGP;
                        }
                        if (a.logp_all) {
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                int nn = lane * 4 + k;
                                if (nn < N) a.logp_all[((size_t)r * T + t) * N + nn] = (xv[k] - m) - lse;
                            }
                        }
                    }
                }
            }
            // feasibility of the chosen action (utils/decoding.py:377-379,393-395)
            {
                const unsigned fb = __shfl_sync(kFull, feas, act >> 2);
                if (!((fb >> (act & 3)) & 1u)) status |= RL4CO_ST_INFEASIBLE_ACTION;
            }
            if (tid == 0) s_act[t] = act;
            // ---- transition
            const unsigned abit = (lane == (act >> 2)) ? (1u << (act & 3)) : 0u;
            if (ENV == RL4CO_ENV_TSP) {                 // tsp/env.py:67-93
                feas &= ~abit;
                if (t == 0) {
                    // from step 1 on: ctx = Wc [h_first ; h_cur] (context.py:123-131).  The
                    // first-node half is constant, fold it into the constant scores.
                    __syncthreads();
                    if (tid < E)
                        sqb[tid] = (a.gctx ? a.gctx[b * E + tid] : 0.f) +
                                   a.tables[((size_t)4 * M + b * (int64_t)N + act) * E + tid];
                    __syncthreads();
                    dot_k(sqb, sb0, sb1);
                    use_c = true;
                }
            } else if (ENV == RL4CO_ENV_CVRP) {         // cvrp/env.py:73-105,137-149
                const float sel = s_dem[act == 0 ? 1 : act];
                used = __fmul_rn(__fadd_rn(used, sel), act != 0 ? 1.f : 0.f);
                vis |= abit;
                feas = cvrp_mask(act);
                done = __all_sync(kFull, vis == own);
            } else {                                    // slap/env.py:38-93
                feas &= ~abit;
            }
            cur = act;
            ++t;
        }

        // ================================================== epilogue =====================
        __syncthreads();
        const int nact = t;
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
        float llp = 0.f;
        if (warp == 1) {
            for (int q = lane; q < nact; q += 32) llp += s_lp[q];
            llp = warp_sum(llp);
        }
        if (lane == 0) spart[warp] = part;
        if (warp == 1 && lane == 0) spart[32] = llp;
        status = __reduce_or_sync(kFull, status);
        if (lane == 0) spart[40 + warp] = __int_as_float(status);
        __syncthreads();
        if (tid == 0) {
            float len = 0.f;
            int st = 0;
            for (int g = 0; g < kWarps; ++g) { len += spart[g]; st |= __float_as_int(spart[40 + g]); }
            a.reward[r] = -len;
            a.ll[r] = spart[32];
            if (a.steps) a.steps[r] = nact;
            if (st) raise_status(a.status, st);
        }
        for (int q = tid; q < T; q += kThreads) {    // pad with depot / log-prob 0 after done
            a.actions[(size_t)r * T + q] = q < nact ? (int64_t)s_act[q] : 0;
            if (a.logp) a.logp[(size_t)r * T + q] = q < nact ? s_lp[q] : 0.f;
        }
        if (a.logp_all)
            for (size_t k = (size_t)nact * N + tid; k < (size_t)T * N; k += kThreads)
                a.logp_all[(size_t)r * T * N + k] = 0.f;
        if (ENV == RL4CO_ENV_SLAP && a.assignment)
            for (int q = tid; q < a.P; q += kThreads) a.assignment[(size_t)r * a.P + q] = q < nact ? s_act[q] : -1;
    }
}

template <int ENV, int G>
static int launch(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    const Layout L = make_layout(a.N, a.T, G, ENV != RL4CO_ENV_SLAP);
    if (L.total > (size_t)227 * 1024) {
        set_error("am_rollout(rf): N=%d T=%d needs %zu B shared memory", a.N, a.T, L.total);
        return 1;
    }
    cudaError_t e = cudaFuncSetAttribute(am_rollout_rf_kernel<ENV, G>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
    if (e != cudaSuccess) {
        set_error("am_rollout(rf): cannot reserve %zu B of shared memory: %s", L.total, cudaGetErrorString(e));
        return 2;
    }
    const int64_t grid = a.B < sm_count ? a.B : sm_count;   // persistent: one CTA per SM
    am_rollout_rf_kernel<ENV, G><<<(unsigned)grid, kThreads, L.total, stream>>>(a);
    RL4CO_CHECK_LAUNCH("am_rollout(rf)");
    return 0;
}

template <int ENV>
static int launch_env(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    const int G = (a.N + kWarps - 1) / kWarps;
    if (G <= 2) return launch<ENV, 2>(a, stream);
    if (G <= 4) return launch<ENV, 4>(a, stream);
    if (G <= 6) return launch<ENV, 6>(a, stream);
    if (G == 7) return launch<ENV, 7>(a, stream);
    return launch<ENV, 8>(a, stream);
}

}  // namespace rf

bool rollout_rf_supports(const rl4co_am_rollout_args &a) {
    if (a.S > 1 || a.multistart || a.N > rf::kSlots) return false;
    const int G = (a.N + rf::kWarps - 1) / rf::kWarps;
    const int Gt = G <= 2 ? 2 : G <= 4 ? 4 : G <= 6 ? 6 : G == 7 ? 7 : 8;
    return rf::make_layout(a.N, a.T, Gt, a.env != RL4CO_ENV_SLAP).total <= (size_t)227 * 1024;
}

int launch_rollout_rf(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    switch (a.env) {
        case RL4CO_ENV_TSP: return rf::launch_env<RL4CO_ENV_TSP>(a, stream);
        case RL4CO_ENV_CVRP: return rf::launch_env<RL4CO_ENV_CVRP>(a, stream);
        default: return rf::launch_env<RL4CO_ENV_SLAP>(a, stream);
    }
}

}  // namespace rl4co
