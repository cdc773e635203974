// Backward of the teacher-forced log-likelihood (first half): given the tours, the per-node tables
// and dL/d(ll) per rollout, the gradients with respect to the tables, the graph context, the step-0 context (tsp) and
// the capacity column (cvrp).  REINFORCE needs exactly this (rl4co/models/rl/reinforce/reinforce.py:73-115: the loss is
// a weighted sum of log-likelihoods); the gradients with respect to the encoder output h and the decoder weights then
// follow from the table projection, which the host keeps as differentiable torch ops (training.py).
//
// Recompute based: nothing is saved by the rollout.  One CTA of 256 threads per instance walks its rollouts step by
// step in the ownership of am_rollout_rf.cu (warp = attention head, lane = (half head, node group), a thread owns nodes
// g, g + 16, ... and 8 dims), recomputes the step's forward quantities
//     q -> masked scores s -> e = 2^(s - max), Z_h -> glimpse o_h / Z_h -> logits -> x = clip tanh(.) -> log-softmax
// (rl4co/models/zoo/am/decoder.py:134-200, rl4co/models/nn/attention.py:263-309, rl4co/utils/decoding.py:141-191) and
// runs the chain rule backwards through them:
//     g_x[n]   = w (1[n = a] - softmax(x)[n])                      w = dL/d ll of the rollout
//     g_z[n]   = g_x[n] clip (1 - tanh^2) / sqrt(E)
//     g_L[n]  += g_z[n] o~          g_o~ = sum_n g_z[n] L[n]       o~ = o / Z (per head)
//     g_V[n]  += p~[n] g_o~         g_s[n] = ln2 p~[n] (<g_o~, V[n]> - <g_o~, o~>)     p~ = e / Z
//     g_K[n]  += g_s[n] q_s         g_q = scale sum_n g_s[n] K[n]                      q_s = scale q
//     g_q -> context row of the current node, graph context, first-node row / step-0 context (tsp), capacity column (cvrp)
// The three table gradients of the instance (3 x N x 128 fp32) and the context-row gradients (N x 128) accumulate in
// SHARED MEMORY over all steps and starts of the instance and leave as plain stores at the end; the three tables
// themselves sit in REGISTERS for the whole instance, as in the forward kernel (streaming them from L2 six times per
// step made the first version 3x slower).
#include "rollout_util.cuh"

namespace rl4co {
namespace bwd {

constexpr int E = RL4CO_EMBED_DIM;
constexpr int kThreads = 256, kWarps = 8;
constexpr int PLS = 144;                       // row stride of the partial-logit matrix [head][node]
constexpr float kScoreScale = 0.25f * 1.4426950408889634f;   // 1/sqrt(16) * log2(e): scores live in base 2
constexpr float kInvSqrtE = 0.08838834764831845f;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

struct Layout {
    size_t off_g, off_gc, off_pl, off_so, off_go, off_x, off_th, off_dem, off_acc, total;
};
__host__ __device__ inline Layout make_layout(int N) {
    Layout L;
    size_t o = 0;
    L.off_g = o;   o += (size_t)3 * N * E * 4;      // gradients of tables 0-2 (glimpse K, glimpse V, folded logit K)
    L.off_gc = o;  o += (size_t)N * E * 4;          // gradient of table 3 (context rows)
    L.off_pl = o;  o += (size_t)kWarps * PLS * 4;   // partial logits [head][node]
    L.off_so = o;  o += E * 4;                      // (unused slack kept for alignment of the vectors below)
    L.off_go = o;  o += E * 4;
    L.off_x = o;   o += 128 * 4;                    // processed logits of the step
    L.off_th = o;  o += 128 * 4;                    // tanh of the step's raw logits
    L.off_dem = o; o += 128 * 4;                    // cvrp: demand of node n
    L.off_acc = o; o += 4 * E * 4;                  // accumulators: graph context | first-node row | step-0 context | capacity column
    L.total = o;
    return L;
}

__device__ __forceinline__ float ex2f(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void ld8(const float *p, float (&v)[8]) {       // 8 consecutive floats, 32-byte aligned
    const float4 a = __ldg(reinterpret_cast<const float4 *>(p));
    const float4 b = __ldg(reinterpret_cast<const float4 *>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ float dot8(const float (&a)[8], const float (&b)[8]) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s = fmaf(a[k], b[k], s);
    return s;
}
// acc[0..7] += c * v[0..7] on a 32-byte aligned shared-memory address owned by this thread alone
__device__ __forceinline__ void axpy8_smem(float *acc, float c, const float (&v)[8]) {
    float4 a = *reinterpret_cast<float4 *>(acc), b = *(reinterpret_cast<float4 *>(acc) + 1);
    a.x = fmaf(c, v[0], a.x); a.y = fmaf(c, v[1], a.y); a.z = fmaf(c, v[2], a.z); a.w = fmaf(c, v[3], a.w);
    b.x = fmaf(c, v[4], b.x); b.y = fmaf(c, v[5], b.y); b.z = fmaf(c, v[6], b.z); b.w = fmaf(c, v[7], b.w);
    *reinterpret_cast<float4 *>(acc) = a;
    *(reinterpret_cast<float4 *>(acc) + 1) = b;
}
// sum over the 16 node groups of a half head (lanes with the same lane >> 4)
__device__ __forceinline__ void allreduce16(float (&v)[8]) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1)
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] += __shfl_xor_sync(kFull, v[k], o);
}
__device__ __forceinline__ float allreduce16(float v) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

template <int ENV, int M>
__global__ void __launch_bounds__(kThreads, 1)
am_evaluate_backward_kernel(const rl4co_am_rollout_args a, const float *__restrict__ grad_ll, float *__restrict__ g_tables,
                            float *__restrict__ g_gctx, float *__restrict__ g_q0, float *__restrict__ g_wcap, const int n_parts) {
    constexpr bool TSP = ENV == RL4CO_ENV_TSP, CVRP = ENV == RL4CO_ENV_CVRP;
    extern __shared__ __align__(128) unsigned char smem[];
    const int N = a.N, T = a.T;
    const Layout L = make_layout(N);
    float *gacc = reinterpret_cast<float *>(smem + L.off_g);       // [3][N][E]
    float *gctab = reinterpret_cast<float *>(smem + L.off_gc);     // [N][E]
    float *spl = reinterpret_cast<float *>(smem + L.off_pl);
    float *sx = reinterpret_cast<float *>(smem + L.off_x);
    float *sth = reinterpret_cast<float *>(smem + L.off_th);
    float *s_dem = reinterpret_cast<float *>(smem + L.off_dem);
    float *acc_gctx = reinterpret_cast<float *>(smem + L.off_acc);
    float *acc_first = acc_gctx + E, *acc_q0 = acc_gctx + 2 * E, *acc_wcap = acc_gctx + 3 * E;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int hh = lane >> 4, g = lane & 15;
    const int dim0 = 16 * warp + 8 * hh;            // first of the 8 dims this thread owns
    const int64_t Mrows = a.B * (int64_t)N;
    const float clip = a.tanh_clipping;
    const int S = a.S > 1 ? a.S : 1;

    if (tid < E) acc_q0[tid] = acc_wcap[tid] = 0.f;     // summed over every instance of this CTA, flushed once
    // work item = (instance, range of its starts): with fewer instances than SMs the starts of an instance are spread
    // over n_parts CTAs, whose partial table gradients then meet in global memory through atomics
    const int s_per = (S + n_parts - 1) / n_parts;
    const int64_t W = a.B * (int64_t)n_parts;
    for (int64_t wi = blockIdx.x; wi < W; wi += gridDim.x) {
        const int64_t b = wi / n_parts;
        const int s_begin = (int)(wi % n_parts) * s_per, s_end = s_begin + s_per < S ? s_begin + s_per : S;
        __syncthreads();                            // the previous instance's flush is done with shared memory
        for (size_t i = tid; i < (size_t)4 * N * E / 4; i += kThreads) reinterpret_cast<float4 *>(gacc)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tid < E) acc_gctx[tid] = 0.f;
        if (tid < 128) s_dem[tid] = (CVRP && tid >= 1 && tid < N) ? a.demand[b * (int64_t)(N - 1) + tid - 1] : 0.f;
        __syncthreads();
        const float *tabK = a.tables + ((size_t)0 * Mrows + b * (int64_t)N) * E + dim0;
        const float *tabV = a.tables + ((size_t)1 * Mrows + b * (int64_t)N) * E + dim0;
        const float *tabL = a.tables + ((size_t)2 * Mrows + b * (int64_t)N) * E + dim0;
        const float *tabC = a.tables + ((size_t)3 * Mrows + b * (int64_t)N) * E + dim0;
        const float *tabF = TSP ? a.tables + ((size_t)4 * Mrows + b * (int64_t)N) * E + dim0 : nullptr;
        // the instance's three tables in registers for all of its steps and starts: this thread's nodes x its 8 dims
        float Kr[M][8], Vr[M][8], Lr[M][8];
#pragma unroll
        for (int j = 0; j < M; ++j) {
            if ((g + 16 * j) < N) {
                ld8(tabK + (size_t)(g + 16 * j) * E, Kr[j]);
                ld8(tabV + (size_t)(g + 16 * j) * E, Vr[j]);
                ld8(tabL + (size_t)(g + 16 * j) * E, Lr[j]);
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) Kr[j][k] = Vr[j][k] = Lr[j][k] = 0.f;
            }
        }
        float gcx[8];                               // graph context at this thread's dims
#pragma unroll
        for (int k = 0; k < 8; ++k) gcx[k] = 0.f;
        if (a.gctx) ld8(a.gctx + b * E + dim0, gcx);
        float wc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) wc[k] = 0.f;
        if (CVRP) ld8(a.wcap + dim0, wc);
        const float cap = CVRP ? a.vehicle_capacity[b] : 0.f;
        unsigned own = 0;
#pragma unroll
        for (int j = 0; j < M; ++j) own |= (g + 16 * j < N) ? (1u << j) : 0u;

        for (int s = s_begin; s < s_end; ++s) {
            const int64_t r = (int64_t)s * a.B + b;             // rollout, start-major (utils/ops.py:11-34)
            const float w = grad_ll[r];
            const int64_t *acts = a.actions_in + (size_t)r * T;
            if (tid < E) acc_first[tid] = 0.f;
            __syncthreads();
            // ---- environment state replayed from the actions (tsp/env.py:67-93, cvrp/env.py:73-149), owner layout
            unsigned alive = own, vis = 0;
            float used = 0.f;
            int cur = 0, first = 0, t = 0;
            bool done = false;
            auto cvrp_mask = [&](int at) {
                unsigned f = 0;
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    const bool exceeds = __fadd_rn(s_dem[g + 16 * j], used) > cap;
                    if (!exceeds && !((vis >> j) & 1u)) f |= 1u << j;
                }
                f &= own;
                if (g == 0) f &= ~1u;
                const bool depot = !(at == 0 && __any_sync(kFull, f != 0));
                if (g == 0 && depot) f |= 1u;
                alive = f;
            };
            auto apply = [&](int act) {
                const unsigned abit = (g == (act & 15)) ? (1u << (act >> 4)) : 0u;
                if (TSP) {
                    alive &= ~abit;
                    if (t == 0) first = act;
                    done = t + 1 >= N;
                } else {
                    const float sel = s_dem[act == 0 ? 1 : act];
                    used = __fmul_rn(__fadd_rn(used, sel), act != 0 ? 1.f : 0.f);
                    vis |= abit;
                    cvrp_mask(act);
                    done = __all_sync(kFull, vis == own);
                }
                cur = act;
            };
            if (CVRP) cvrp_mask(0);
            if (a.multistart) {                     // forced POMO start node, log-prob 0: no gradient (decoding.py:287-305)
                apply((int)acts[0]);
                t = 1;
            }
            const int t_end = TSP ? (N < T ? N : T) : T;
            if (w == 0.f) t = t_end;                // nothing flows into this rollout

            while (!done && t < t_end) {
                const int act = (int)acts[t];
                // ---- query (am/decoder.py:141-142), pre-scaled: q_s = scale (gctx + ctx [+ capacity column * remaining])
                float q[8];
                const bool step0 = TSP && t == 0;   // single start only: a multistart rollout begins at t = 1
                if (step0) {
                    ld8(a.q0 + dim0, q);
                } else {
                    ld8(tabC + (size_t)cur * E, q);
                    if (TSP) {
                        float f8[8];
                        ld8(tabF + (size_t)first * E, f8);
#pragma unroll
                        for (int k = 0; k < 8; ++k) q[k] += f8[k];
                    }
                }
                const float rem = CVRP ? __fsub_rn(cap, used) : 0.f;
#pragma unroll
                for (int k = 0; k < 8; ++k) q[k] = (q[k] + gcx[k] + (CVRP ? wc[k] * rem : 0.f)) * kScoreScale;

                // ================================================== forward of the step, recomputed
                float e[M], mx = -INFINITY;
#pragma unroll
                for (int j = 0; j < M; ++j) {       // masked glimpse scores (attention.py:289-306), base 2
                    float d = dot8(Kr[j], q);           // rows of nodes >= N are zero
                    d += __shfl_xor_sync(kFull, d, 16);
                    e[j] = ((alive >> j) & 1u) ? d : -INFINITY;
                    mx = fmaxf(mx, e[j]);
                }
                mx = warp_max(mx);
                float zsum = 0.f;
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    e[j] = ex2f(e[j] - mx);         // masked: 2^-inf = 0
                    zsum += e[j];
                }
                zsum = allreduce16(zsum);
                const float invz = 1.f / zsum;
                float o[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = 0.f;
#pragma unroll
                for (int j = 0; j < M; ++j)
#pragma unroll
                    for (int k = 0; k < 8; ++k) o[k] = fmaf(e[j], Vr[j][k], o[k]);      // masked: e = 0
                allreduce16(o);
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] *= invz;           // o~: normalised glimpse, this thread's 8 dims
#pragma unroll
                for (int j = 0; j < M; ++j) {       // partial pointer logits of this head (project_out folded into table 2)
                    float d = dot8(Lr[j], o);
                    d += __shfl_xor_sync(kFull, d, 16);
                    if (hh == 0 && ((own >> j) & 1u)) spl[warp * PLS + g + 16 * j] = d;
                }
                __syncthreads();
                if (tid < 128) {                    // thread = node tid = g + 16 (2 warp + hh): its feasibility bit is local
                    const int jb = 2 * warp + hh;
                    float xv = -INFINITY, th = 0.f;
                    if (jb < M && ((own >> jb) & 1u)) {
                        const float *col = spl + tid;
                        const float sum = ((col[0] + col[2 * PLS]) + (col[4 * PLS] + col[6 * PLS])) +
                                          ((col[PLS] + col[3 * PLS]) + (col[5 * PLS] + col[7 * PLS]));
                        const float e2 = ex2f(sum * (kInvSqrtE * 2.f * kLog2e));
                        th = 1.f - __fdividef(2.f, e2 + 1.f);       // tanh(sum / sqrt(E)) (decoding.py:172-173)
                        if ((alive >> jb) & 1u) xv = clip * th;
                    }
                    sx[tid] = xv;
                    sth[tid] = th;
                }
                __syncthreads();
                // log-softmax ingredients of the row (every warp, redundantly)
                float4 xr = *reinterpret_cast<const float4 *>(sx + 4 * lane);
                const float m = warp_max(fmaxf(fmaxf(xr.x, xr.y), fmaxf(xr.z, xr.w)));
                const float zz = warp_sum((ex2f((xr.x - m) * kLog2e) + ex2f((xr.y - m) * kLog2e)) +
                                          (ex2f((xr.z - m) * kLog2e) + ex2f((xr.w - m) * kLog2e)));
                const float inv_zz = 1.f / zz;

                // ================================================== backward of the step
                // d ll / d raw logit of this thread's nodes: w (1[n = a] - softmax) clip (1 - tanh^2) / sqrt(E)
                float gz[M];
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    const int n = g + 16 * j;
                    float v = 0.f;
                    if ((alive >> j) & 1u) {
                        const float pi = ex2f((sx[n] - m) * kLog2e) * inv_zz;
                        const float th = sth[n];
                        v = w * ((n == act ? 1.f : 0.f) - pi) * clip * (1.f - th * th) * kInvSqrtE;
                    }
                    gz[j] = v;
                }
                // logits = sum over heads of <o~, L[n]>:  g_L[n] += g_z[n] o~,  g_o~ = sum_n g_z[n] L[n]
                float go[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) go[k] = 0.f;
#pragma unroll
                for (int j = 0; j < M; ++j)
                    if ((alive >> j) & 1u) {
                        const int n = g + 16 * j;
#pragma unroll
                        for (int k = 0; k < 8; ++k) go[k] = fmaf(gz[j], Lr[j][k], go[k]);
                        axpy8_smem(gacc + ((size_t)2 * N + n) * E + dim0, gz[j], o);
                    }
                allreduce16(go);
                // o~ = sum_n p~[n] V[n], p~ = softmax(s):  g_V[n] += p~[n] g_o~,  g_s[n] = ln2 p~[n] (<g_o~, V[n]> - <g_o~, o~>)
                float dotgo = dot8(go, o);
                dotgo += __shfl_xor_sync(kFull, dotgo, 16);
                float gs[M];
#pragma unroll
                for (int j = 0; j < M; ++j) {
                    const int n = g + 16 * j;
                    const bool al = (alive >> j) & 1u;
                    const float pt = e[j] * invz;
                    float gp = 0.f;
                    if (al) {
                        gp = dot8(go, Vr[j]);
                        axpy8_smem(gacc + ((size_t)1 * N + n) * E + dim0, pt, go);
                    }
                    gp += __shfl_xor_sync(kFull, gp, 16);
                    gs[j] = al ? kLn2 * pt * (gp - dotgo) : 0.f;
                }
                // s[n] = <K[n], q_s>:  g_K[n] += g_s[n] q_s,  g_q = scale sum_n g_s[n] K[n]
                float gq[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) gq[k] = 0.f;
#pragma unroll
                for (int j = 0; j < M; ++j)
                    if ((alive >> j) & 1u) {
                        const int n = g + 16 * j;
#pragma unroll
                        for (int k = 0; k < 8; ++k) gq[k] = fmaf(gs[j], Kr[j][k], gq[k]);
                        axpy8_smem(gacc + ((size_t)0 * N + n) * E + dim0, gs[j], q);
                    }
                allreduce16(gq);
                if (g == 0) {                       // one lane per half head carries its 8 dims of g_q to where q came from
#pragma unroll
                    for (int k = 0; k < 8; ++k) gq[k] *= kScoreScale;
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        acc_gctx[dim0 + k] += gq[k];
                        if (step0) acc_q0[dim0 + k] += gq[k];
                        else {
                            gctab[(size_t)cur * E + dim0 + k] += gq[k];
                            if (TSP) acc_first[dim0 + k] += gq[k];
                        }
                        if (CVRP) acc_wcap[dim0 + k] = fmaf(rem, gq[k], acc_wcap[dim0 + k]);
                    }
                }
                apply(act);
                ++t;
            }
            __syncthreads();
            if (TSP && tid < E && w != 0.f) {       // the first node's half of the context: one row per rollout
                float *row = g_tables + ((size_t)4 * Mrows + b * (int64_t)N + first) * E;
                if (n_parts > 1) atomicAdd(row + tid, acc_first[tid]);
                else row[tid] += acc_first[tid];
            }
        }
        // ---- the instance's table gradients leave shared memory (this CTA owns every row of the instance)
        __syncthreads();
        if (n_parts == 1) {                         // this CTA owns every row of the instance: plain stores
            for (size_t i = tid; i < (size_t)N * E / 4; i += kThreads) {
#pragma unroll
                for (int tb = 0; tb < 3; ++tb)
                    reinterpret_cast<float4 *>(g_tables + ((size_t)tb * Mrows + b * (int64_t)N) * E)[i] =
                        reinterpret_cast<const float4 *>(gacc + (size_t)tb * N * E)[i];
                reinterpret_cast<float4 *>(g_tables + ((size_t)3 * Mrows + b * (int64_t)N) * E)[i] = reinterpret_cast<const float4 *>(gctab)[i];
            }
            if (g_gctx && tid < E) g_gctx[b * E + tid] = acc_gctx[tid];
        } else {                                    // several CTAs share the instance: accumulate (buffers are zeroed by the caller)
            for (size_t i = tid; i < (size_t)N * E; i += kThreads) {
#pragma unroll
                for (int tb = 0; tb < 3; ++tb)
                    atomicAdd(g_tables + ((size_t)tb * Mrows + b * (int64_t)N) * E + i, gacc[(size_t)tb * N * E + i]);
                atomicAdd(g_tables + ((size_t)3 * Mrows + b * (int64_t)N) * E + i, gctab[i]);
            }
            if (g_gctx && tid < E) atomicAdd(g_gctx + b * E + tid, acc_gctx[tid]);
        }
    }
    __syncthreads();
    if (tid < E) {
        if (g_q0 && TSP) atomicAdd(g_q0 + tid, acc_q0[tid]);
        if (g_wcap && CVRP) atomicAdd(g_wcap + tid, acc_wcap[tid]);
    }
}

template <int ENV, int M>
static int launch(const rl4co_am_rollout_args &a, const float *grad_ll, float *g_tables, float *g_gctx, float *g_q0,
                  float *g_wcap, cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    const Layout L = make_layout(a.N);
    cudaError_t e = cudaFuncSetAttribute(am_evaluate_backward_kernel<ENV, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
    if (e != cudaSuccess) {
        set_error("am_evaluate_backward: cannot reserve %zu B of shared memory: %s", L.total, cudaGetErrorString(e));
        return 2;
    }
    const int S = a.S > 1 ? a.S : 1;
    int64_t n_parts = 1;
    if (S > 1 && a.B < sm_count) {
        n_parts = (sm_count + a.B - 1) / a.B;
        if (n_parts > S) n_parts = S;
        const int64_t per = (S + n_parts - 1) / n_parts;
        n_parts = (S + per - 1) / per;              // no empty parts
    }
    const int64_t W = a.B * n_parts;
    const int64_t grid = W < sm_count ? W : sm_count;
    am_evaluate_backward_kernel<ENV, M><<<(unsigned)grid, kThreads, L.total, stream>>>(a, grad_ll, g_tables, g_gctx, g_q0, g_wcap,
                                                                                      (int)n_parts);
    RL4CO_CHECK_LAUNCH("am_evaluate_backward");
    return 0;
}

template <int ENV>
static int launch_env(const rl4co_am_rollout_args &a, const float *grad_ll, float *g_tables, float *g_gctx, float *g_q0,
                      float *g_wcap, cudaStream_t stream) {
    const int m = a.N <= 32 ? 2 : a.N <= 64 ? 4 : a.N <= 112 ? 7 : 8;
    switch (m) {
        case 2: return launch<ENV, 2>(a, grad_ll, g_tables, g_gctx, g_q0, g_wcap, stream);
        case 4: return launch<ENV, 4>(a, grad_ll, g_tables, g_gctx, g_q0, g_wcap, stream);
        default: return launch<ENV, 7>(a, grad_ll, g_tables, g_gctx, g_q0, g_wcap, stream);
    }
}

}  // namespace bwd
}  // namespace rl4co

using namespace rl4co;

extern "C" int rl4co_am_evaluate_backward_max_nodes(void) {
    int n = 128;
    while (n > 2 && bwd::make_layout(n).total > (size_t)227 * 1024) --n;
    return n;
}

extern "C" int rl4co_am_evaluate_backward(const rl4co_am_rollout_args *args, const float *grad_ll, float *g_tables,
                                          float *g_gctx, float *g_q0, float *g_wcap, rl4co_stream_t stream) {
    RL4CO_REQUIRE(args != nullptr, "am_evaluate_backward: null args");
    const rl4co_am_rollout_args &a = *args;
    if (a.B == 0) return 0;
    RL4CO_REQUIRE(a.env == RL4CO_ENV_TSP || a.env == RL4CO_ENV_CVRP, "am_evaluate_backward: tsp and cvrp only, got env %d", a.env);
    RL4CO_REQUIRE(a.B > 0 && a.N >= 2 && a.T >= 1, "am_evaluate_backward: bad sizes B=%lld N=%d T=%d", (long long)a.B, a.N, a.T);
    RL4CO_REQUIRE(a.N <= rl4co_am_evaluate_backward_max_nodes(),
                  "am_evaluate_backward: N=%d exceeds the %d nodes whose gradient accumulators fit in shared memory", a.N,
                  rl4co_am_evaluate_backward_max_nodes());
    RL4CO_REQUIRE(a.tables && a.actions_in && grad_ll && g_tables, "am_evaluate_backward: null tensor");
    RL4CO_REQUIRE(a.tanh_clipping > 0.f && a.temperature == 1.f && a.mask_logits != 0,
                  "am_evaluate_backward: standard logit post-processing only (tanh clipping > 0, temperature 1, masked logits)");
    RL4CO_REQUIRE(a.n_tables >= (a.env == RL4CO_ENV_TSP ? 5 : 4), "am_evaluate_backward: missing tables");
    if (a.env == RL4CO_ENV_TSP) {
        RL4CO_REQUIRE(a.q0 != nullptr, "am_evaluate_backward: tsp needs q0");
        return bwd::launch_env<RL4CO_ENV_TSP>(a, grad_ll, g_tables, g_gctx, g_q0, g_wcap, (cudaStream_t)stream);
    }
    RL4CO_REQUIRE(a.wcap && a.demand && a.vehicle_capacity, "am_evaluate_backward: cvrp needs wcap / demand / capacity");
    RL4CO_REQUIRE(!a.multistart || a.num_loc >= 1, "am_evaluate_backward: bad multistart setup");
    return bwd::launch_env<RL4CO_ENV_CVRP>(a, grad_ll, g_tables, g_gctx, g_q0, g_wcap, (cudaStream_t)stream);
}
