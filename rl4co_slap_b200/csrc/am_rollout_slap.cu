// K2 (SLAP specialisation) - persistent Attention-Model rollout for the storage-location-assignment env.
//
// What SLAP allows that the routing envs do not: its context embedding returns zeros (`SLAPContext`,
// examples/slap.py:11-84 - the wiring BASELINE.json names), so the decoder query is the projected graph context
// alone (am/decoder.py:141-142) and the glimpse QUERY IS THE SAME AT EVERY STEP
// of a trajectory and so are the glimpse scores s[h][n] = <q_h, K[n]_h>.  Only the mask changes, and it only
// ever loses the one node that was just assigned (slap/env.py:38-93).  The per-step glimpse (attention.py:289-306)
//     o_h = sum_{n alive} 2^(s[n] - m) V[n]_h / sum_{n alive} 2^(s[n] - m),      m = max_{n alive} s[n]
// is therefore maintained INCREMENTALLY: numerator O_h and denominator Z_h are kept in fp64 and lose the term of
// the assigned node, O_h -= e_a V[a]_h, Z_h -= e_a.  As long as the node that holds the maximum is alive, m is the
// value the reference's softmax subtracts too, every e_n is bit-identical to the one the step-by-step kernel
// forms, Z >= 1, and a subtraction cancels at most half of the magnitude (fp64 keeps the accumulated error
// ~1e-15).  When the maximum node itself is assigned (its e is exactly 1.0f) the head is recomputed from its
// stored scores with the new maximum - one pass over the V rows, from L2.
//
// Per step that leaves ONE mat-vec, the pointer logits <o, L'[n]> (project_out folded into L', attention.py:
// 276-282), instead of three; only L' stays in registers (56 per thread at N <= 112), so TWO CTAs fit an SM and
// cover each other's barrier and shuffle latency.  K and V stream through once per instance.
//
// Ownership as in am_rollout_rf.cu: warp = head, lane = (hh, g), the thread owns nodes g + 16 j and 8 dims.
// Selection, Gumbel noise keying, log-prob bookkeeping and outputs are the same as in that kernel, so for the
// same seed a sampled rollout picks the same nodes unless a logit differs in the last bits.
//
// Reference op order being reproduced: SURVEY.md appendix A; rl4co/models/zoo/am/decoder.py:134-200,
// rl4co/models/nn/attention.py:263-309, rl4co/utils/decoding.py:141-191,327-397,
// rl4co/envs/warehousing/slap/env.py:38-143.
#include "rollout_util.cuh"

namespace rl4co {
namespace slapk {

typedef unsigned long long u64;

constexpr int E = RL4CO_EMBED_DIM;   // 128
constexpr int kThreads = 256;
constexpr int kWarps = 8;
constexpr float kScoreScale = 0.25f * 1.4426950408889634f;   // 1/sqrt(16) * log2(e)
constexpr float kInvSqrtE = 0.08838834764831845f;             // 1/sqrt(128)
constexpr float kLog2e = 1.4426950408889634f;
constexpr int PLS = 144;              // row stride of the partial-logit matrix [head][node]
constexpr size_t kSmemLimit = 112 * 1024;   // two CTAs per SM

struct Layout {
    size_t off_pl, off_xs, off_sc, off_e, off_o, off_mw, off_act, off_ain, off_lp, off_locs, off_noise, off_part, off_pick, off_bar, off_v, total;
};
__host__ __device__ inline size_t up16(size_t v) { return (v + 15) & ~(size_t)15; }
__host__ __device__ inline Layout make_layout(int N, int P, int M, int n_pick) {
    Layout L;
    size_t o = 0;
    L.off_pl = o;    o += (size_t)kWarps * PLS * 4;          // partial logits [head][node]
    L.off_xs = o;    o += (size_t)P * 128 * 4;               // processed logits, one row per step
    L.off_sc = o;    o += (size_t)M * kThreads * 4;          // glimpse scores of the own nodes [j][thread]
    L.off_e = o;     o += (size_t)kWarps * 128 * 4;          // e[head][node] = 2^(s - m), 0 once assigned
    L.off_o = o;     o += E * 4;                             // glimpse numerators after a recompute (warp private)
    L.off_mw = o;    o += up16((size_t)P * 4);
    L.off_act = o;   o += up16((size_t)P * 4);
    L.off_ain = o;   o += up16((size_t)P * 4);
    L.off_lp = o;    o += up16((size_t)P * 4);
    L.off_locs = o;  o += up16((size_t)N * 8);
    L.off_noise = o; o += 16 * 128 * 4;                      // sampling: Gumbel noise of 16 steps
    L.off_part = o;  o += 64 * 4;
    L.off_pick = o;  o += up16((size_t)n_pick * 4);          // picklist as int
    L.off_bar = o;   o += 16;
    o = (o + 127) & ~(size_t)127;
    L.off_v = o;     o += (size_t)N * E * 4;                 // glimpse values [node][128], landed by TMA
    L.total = o;
    return L;
}

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
__device__ __forceinline__ float fast_tanh(float x) {
    const float t = ex2f(x * 2.8853900817779268f);
    return 1.f - __fdividef(2.f, t + 1.f);
}
__device__ __forceinline__ float redux_max(float v) {
    float r;
    asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ unsigned redux_min_u32(unsigned v) {
    unsigned r;
    asm volatile("redux.sync.min.u32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(v));
    return r;
}
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {
    u64 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ u64 pack2(float lo, float hi) {
    u64 d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(lo), "f"(hi));
    return d;
}
__device__ __forceinline__ void unpack2(u64 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ float hsum2(u64 v) {
    float lo, hi;
    unpack2(v, lo, hi);
    return lo + hi;
}
__device__ __forceinline__ void ldg_256(const float *p, u64 (&v)[4]) {
    asm volatile("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(p));
}
__device__ __forceinline__ void prefetch_l2(const void *p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

#ifdef SLAPK_TRACE
// -DSLAPK_TRACE (tools/trace_slap.py): thread 0 of CTA 0 stamps clock64() into a.logp_all, reused as a [P + 1][8] int64
// buffer, during its SECOND instance (the other CTA of the SM is then in flight): six points per step, and in row P the
// phases of the instance
#define SLAPK_TRACE_AT(k) do { if (tid == 0 && blockIdx.x == 0 && it == 1) reinterpret_cast<long long *>(a.logp_all)[t * 8 + (k)] = clock64(); } while (0)
#define SLAPK_TRACE_INST(k) do { if (tid == 0 && blockIdx.x == 0 && it == 1) reinterpret_cast<long long *>(a.logp_all)[P * 8 + (k)] = clock64(); } while (0)
#else
#define SLAPK_TRACE_AT(k) do { } while (0)
#define SLAPK_TRACE_INST(k) do { } while (0)
#endif

// error-free pieces of the compensated (two-float) accumulators: value = hi + lo
__device__ __forceinline__ void two_sum(float a, float b, float &s, float &err) {
    s = __fadd_rn(a, b);
    const float bb = __fsub_rn(s, a);
    err = __fadd_rn(__fsub_rn(a, __fsub_rn(s, bb)), __fsub_rn(b, bb));
}
// (hi, lo) -= e * v, with the product's rounding error recovered by an FMA
__device__ __forceinline__ void sub_product(float &hi, float &lo, float e, float v) {
    const float p = __fmul_rn(e, v);
    const float pe = __fmaf_rn(e, v, -p);
    float s, err;
    two_sum(hi, -p, s, err);
    hi = s;
    lo = __fadd_rn(lo, __fsub_rn(err, pe));
}

template <int M>
__global__ void __launch_bounds__(kThreads, 2)
am_rollout_slap_kernel(const rl4co_am_rollout_args a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int N = a.N, T = a.T, P = a.P < a.T ? a.P : a.T;
    const Layout L = make_layout(N, P, M, a.O * a.K);
    float *spl = reinterpret_cast<float *>(smem + L.off_pl);
    float *sxs = reinterpret_cast<float *>(smem + L.off_xs);
    float *s_sc = reinterpret_cast<float *>(smem + L.off_sc);
    float *s_e = reinterpret_cast<float *>(smem + L.off_e);
    float *so = reinterpret_cast<float *>(smem + L.off_o);
    float *s_mw = reinterpret_cast<float *>(smem + L.off_mw);
    int *s_act = reinterpret_cast<int *>(smem + L.off_act);
    int *s_ain = reinterpret_cast<int *>(smem + L.off_ain);
    float *s_lp = reinterpret_cast<float *>(smem + L.off_lp);
    float2 *s_locs = reinterpret_cast<float2 *>(smem + L.off_locs);
    float4 *s_noise = reinterpret_cast<float4 *>(smem + L.off_noise);
    float *spart = reinterpret_cast<float *>(smem + L.off_part);
    int *s_pick = reinterpret_cast<int *>(smem + L.off_pick);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + L.off_bar);
    float *sV = reinterpret_cast<float *>(smem + L.off_v);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int hh = lane >> 4, g = lane & 15;
    const int dim0 = 16 * warp + 8 * hh;            // first of the 8 dims this thread multiplies
    // the ONE of those dims whose glimpse numerator this lane keeps (where the butterfly below leaves it)
    const int kd = ((g & 8) ? 4 : 0) + ((g & 4) ? 2 : 0) + ((g & 2) ? 1 : 0);
    const int64_t Mrows = a.B * (int64_t)N;         // rows per table
    const size_t tab_bytes = (size_t)N * E * 4;
    const bool std_logits = a.tanh_clipping > 0.f && a.temperature == 1.f && a.mask_logits != 0;
    const float clip = a.tanh_clipping;
    const bool sampling = a.mode == RL4CO_MODE_SAMPLING;
    const int n_pick = a.O * a.K;

    unsigned own = 0, rown = 0;
#pragma unroll
    for (int j = 0; j < M; ++j) own |= (g + 16 * j < N) ? (1u << j) : 0u;
#pragma unroll
    for (int k = 0; k < 4; ++k) rown |= (4 * lane + k < N) ? (1u << k) : 0u;

    if (tid == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    unsigned par = 0u;
    auto stage_values = [&](int64_t b) {            // thread 0: TMA the value table of instance b into shared memory
        mbar_expect_tx(bar, (unsigned)tab_bytes);
        const char *src = reinterpret_cast<const char *>(a.tables + ((size_t)1 * Mrows + b * (int64_t)N) * E);
        char *dst = reinterpret_cast<char *>(sV);
        for (size_t off = 0; off < tab_bytes; off += 16384)
            tma_load_1d(dst + off, src + off, (unsigned)(tab_bytes - off < 16384 ? tab_bytes - off : 16384), bar);
    };
    u64 L2[M][4];                                   // logit keys of the current instance, loaded one instance ahead
    auto load_logit_keys = [&](int64_t b) {
        const float *lb = a.tables + ((size_t)2 * Mrows + b * (int64_t)N) * E + dim0;
#pragma unroll
        for (int j = 0; j < M; ++j) {
            const int n = g + 16 * j;
            if (n < N) ldg_256(lb + (size_t)n * E, L2[j]);
            else {
#pragma unroll
                for (int i = 0; i < 4; ++i) L2[j][i] = 0ull;
            }
        }
    };
    if ((int64_t)blockIdx.x < a.B) {
        if (tid == 0) stage_values(blockIdx.x);
        load_logit_keys(blockIdx.x);
    }

    int it = 0;
    for (int64_t b = blockIdx.x; b < a.B; b += gridDim.x, ++it) {
        const int64_t r = b;
        int t = 0, status = 0;
        SLAPK_TRACE_INST(0);
        const float *kb = a.tables + ((size_t)0 * Mrows + b * (int64_t)N) * E + dim0;
        const float *vb = a.tables + ((size_t)1 * Mrows + b * (int64_t)N) * E + dim0;
        // ------------------------------------------------ the one query of the trajectory, pre-scaled
        u64 q2[4];
        {
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = 0.f;
            if (a.gctx) {
                const float4 g0 = __ldg(reinterpret_cast<const float4 *>(a.gctx + b * E + dim0));
                const float4 g1 = __ldg(reinterpret_cast<const float4 *>(a.gctx + b * E + dim0) + 1);
                v[0] = g0.x; v[1] = g0.y; v[2] = g0.z; v[3] = g0.w; v[4] = g1.x; v[5] = g1.y; v[6] = g1.z; v[7] = g1.w;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) q2[i] = pack2(v[2 * i] * kScoreScale, v[2 * i + 1] * kScoreScale);
        }
        // ------------------------------------------------ glimpse scores, once (attention.py:289-306), base-2 scaled
        float s[M];
        {
            u64 kr[M][4];
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const int n = g + 16 * j;
                if (n < N) ldg_256(kb + (size_t)n * E, kr[j]);
                else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) kr[j][i] = 0ull;
                }
            }
            u64 acc[M];
#pragma unroll
            for (int j = 0; j < M; ++j) acc[j] = 0ull;
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < M; ++j) acc[j] = ffma2(kr[j][i], q2[i], acc[j]);
#pragma unroll
            for (int j = 0; j < M; ++j) s[j] = hsum2(acc[j]);
#pragma unroll
            for (int j = 0; j < M; ++j) s[j] += __shfl_xor_sync(kFull, s[j], 16);
        }
        SLAPK_TRACE_INST(1);
        // small per-instance state -> shared memory (the previous instance's epilogue ended with a barrier)
#pragma unroll
        for (int j = 0; j < M; ++j) s_sc[j * kThreads + tid] = s[j];
        {
            const float2 *lrow = reinterpret_cast<const float2 *>(a.locs + b * (int64_t)N * 2);
            for (int n = tid; n < N; n += kThreads) s_locs[n] = lrow[n];
            const int64_t *pl = a.picklist + b * (int64_t)n_pick;
            for (int q = tid; q < n_pick; q += kThreads) s_pick[q] = (int)pl[q];
        }
        if (tid == 0 && b + gridDim.x < a.B) {      // next instance: its three tables -> L2
            const int64_t nb = b + gridDim.x;
            for (int tb = 0; tb < 3; ++tb) {
                const char *src = reinterpret_cast<const char *>(a.tables + ((size_t)tb * Mrows + nb * (int64_t)N) * E);
                for (size_t off = 0; off < tab_bytes; off += 32768)
                    prefetch_l2(src + off, (unsigned)(tab_bytes - off < 32768 ? tab_bytes - off : 32768));
            }
        }
        auto gen_noise = [&](int step) {
            float Gn[4];
            gumbel4(a.seed, a.offset, (uint64_t)r, (uint32_t)step, lane, Gn);
            s_noise[(step & 15) * 32 + lane] = make_float4(Gn[0], Gn[1], Gn[2], Gn[3]);
        };
        if (sampling) {                             // Gumbel noise of steps 0..15 (warp w: steps w and w + 8)
            gen_noise(warp);
            gen_noise(warp + 8);
        }
        if (a.mode == RL4CO_MODE_EVALUATE) {
            for (int q = tid; q < P; q += kThreads) {
                int64_t gv = a.actions_in[(size_t)r * T + q];
                s_ain[q] = (gv < 0 || gv >= N) ? -1 : (int)gv;
            }
        }

        // ------------------------------------------------ env state (slap/env.py:95-129): node 0 is never assignable
        unsigned alive = own, ralive = rown;
        if (g == 0) alive &= ~1u;
        if (lane == 0) ralive &= ~1u;

        // ------------------------------------------------ glimpse state of this head.  Numerator of dim dim0 + kd and the
        // denominator as two-float sums (hi + lo); zref = denominator when the exponentials were last formed
        float Ohi, Olo, Zhi, Zlo, zref;
        // butterfly over the 16 node groups, same order as am_rollout_rf.cu: leaves dim kd of the half head in this lane
        auto reduce_head = [&](const u64 (&oacc)[4], float psum) {
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
            Ohi = w1; Olo = 0.f;
            Zhi = psum; Zlo = 0.f; zref = psum;
            if ((g & 1) == 0) so[dim0 + kd] = w1;
        };
        {   // first glimpse state: the value rows pass through registers once (shared memory gets its own copy by TMA)
            float mx = -INFINITY;
#pragma unroll
            for (int j = 0; j < M; ++j) {
                s[j] = ((alive >> j) & 1u) ? s[j] : -INFINITY;
                mx = fmaxf(mx, s[j]);
            }
            mx = redux_max(mx);
            u64 vr[M][4];
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const int n = g + 16 * j;
                if (n < N) ldg_256(vb + (size_t)n * E, vr[j]);
                else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) vr[j][i] = 0ull;
                }
            }
            float psum = 0.f;
            u64 oacc[4] = {0ull, 0ull, 0ull, 0ull};
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const float p = ex2f(s[j] - mx);        // node 0 / absent: 2^-inf = 0
                psum += p;
                if (hh == 0) s_e[warp * 128 + g + 16 * j] = p;
                const u64 pp = pack2(p, p);
#pragma unroll
                for (int i = 0; i < 4; ++i) oacc[i] = ffma2(vr[j][i], pp, oacc[i]);
            }
            reduce_head(oacc, psum);
        }
        // the same from shared memory, with the maximum of the nodes that are still alive (rare, see the header)
        auto recompute = [&]() {
            float sj[M], mx = -INFINITY;
#pragma unroll
            for (int j = 0; j < M; ++j) {
                sj[j] = ((alive >> j) & 1u) ? s_sc[j * kThreads + tid] : -INFINITY;
                mx = fmaxf(mx, sj[j]);
            }
            mx = redux_max(mx);
            float psum = 0.f;
            u64 oacc[4] = {0ull, 0ull, 0ull, 0ull};
#pragma unroll
            for (int j = 0; j < M; ++j) {
                const float p = ex2f(sj[j] - mx);
                psum += p;
                if (hh == 0) s_e[warp * 128 + g + 16 * j] = p;
                const u64 pp = pack2(p, p);
                const int n = g + 16 * j < N ? g + 16 * j : N - 1;      // absent nodes: p = 0, any row
                const ulonglong2 v0 = *reinterpret_cast<const ulonglong2 *>(sV + (size_t)n * E + dim0);
                const ulonglong2 v1 = *(reinterpret_cast<const ulonglong2 *>(sV + (size_t)n * E + dim0) + 1);
                oacc[0] = ffma2(v0.x, pp, oacc[0]);
                oacc[1] = ffma2(v0.y, pp, oacc[1]);
                oacc[2] = ffma2(v1.x, pp, oacc[2]);
                oacc[3] = ffma2(v1.y, pp, oacc[3]);
            }
            reduce_head(oacc, psum);
        };
        SLAPK_TRACE_INST(2);
        mbar_wait(bar, par);                        // value table landed in shared memory
        par ^= 1u;
        __syncthreads();                            // noise / given actions visible to every warp
        SLAPK_TRACE_INST(3);

        // ================================================== decode loop ==================
        while (t < P) {
            SLAPK_TRACE_AT(0);
            float G[4] = {0.f, 0.f, 0.f, 0.f};
            if (sampling) {                             // noise ring as in am_rollout_rf.cu
                if ((t & 7) == 0 && t > 0) gen_noise(t + 8 + warp);
                const float4 g4 = s_noise[(t & 15) * 32 + lane];
                G[0] = g4.x; G[1] = g4.y; G[2] = g4.z; G[3] = g4.w;
            }
            // ---- partial pointer logits of this head, project_out folded (attention.py:276-282)
            __syncwarp();                               // the glimpse numerators of all 8 dims are in `so`
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
                const float inv = __fdividef(1.f, Zhi + Zlo);     // softmax normaliser of this head
                float pl[M];
#pragma unroll
                for (int j = 0; j < M; ++j) pl[j] = hsum2(acc[j]) * inv;
#pragma unroll
                for (int j = 0; j < M; ++j) pl[j] += __shfl_xor_sync(kFull, pl[j], 16);
                float *row = spl + warp * PLS + g;
                if (hh == 0) {
#pragma unroll
                    for (int j = 0; j < M; ++j) row[16 * j] = pl[j];
                }
            }
            SLAPK_TRACE_AT(1);
            __syncthreads();
            SLAPK_TRACE_AT(2);
            float *sxrow = sxs + t * 128;
            // ---- warps 0-3, thread = node 32 warp + lane = g + 16 (2 warp + hh): logit = sum_h pl_h / sqrt(E),
            //      then process_logits (utils/decoding.py:141-191)
            if (warp < 4) {
                const float *col = spl + tid;
                const float sum = ((col[0] + col[2 * PLS]) + (col[4 * PLS] + col[6 * PLS])) +
                                  ((col[PLS] + col[3 * PLS]) + (col[5 * PLS] + col[7 * PLS]));
                const int jb = 2 * warp + hh;
                const bool exists = (own >> jb) & 1u, feasible = (alive >> jb) & 1u;
                float xv;
                if (std_logits) {
                    const float e2 = ex2f(sum * (kInvSqrtE * 2.f * kLog2e));
                    xv = fmaf(-2.f * clip, rcpf(e2 + 1.f), clip);
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
            SLAPK_TRACE_AT(3);
            __syncthreads();
            SLAPK_TRACE_AT(4);
            // ---- every warp: all logits, lane = nodes 4 lane .. 4 lane + 3
            float x[4];
            {
                const float4 xv = *reinterpret_cast<const float4 *>(sxrow + 4 * lane);
                x[0] = xv.x; x[1] = xv.y; x[2] = xv.z; x[3] = xv.w;
            }
            const float gm = (sampling && std_logits) ? clip : redux_max(fmaxf(fmaxf(x[0], x[1]), fmaxf(x[2], x[3])));
            // ---- selection (utils/decoding.py:371-397,410-499)
            int act;
            if (a.mode == RL4CO_MODE_GREEDY) {
                unsigned cand = 0x7fffffffu;            // argmax, lowest index on ties
                cand = x[3] == gm ? 4u * lane + 3u : cand;
                cand = x[2] == gm ? 4u * lane + 2u : cand;
                cand = x[1] == gm ? 4u * lane + 1u : cand;
                cand = x[0] == gm ? 4u * lane : cand;
                const unsigned win = redux_min_u32(cand);
                act = win < (unsigned)N ? (int)win : 0;
            } else if (a.mode == RL4CO_MODE_EVALUATE) {
                act = s_ain[t];
                if (act < 0) { status |= RL4CO_ST_ACTION_RANGE; act = 0; }
            } else {                                    // Gumbel-max sample (rollout_util.cuh)
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
            if (lane == (act >> 2) && !((ralive >> (act & 3)) & 1u)) status |= RL4CO_ST_INFEASIBLE_ACTION;
            if (tid == 0) { s_act[t] = act; s_mw[t] = gm; }
            // ---- transition (slap/env.py:38-93): the location leaves the mask and this head's glimpse state
            alive &= ~((g == (act & 15)) ? (1u << (act >> 4)) : 0u);
            ralive &= ~((lane == (act >> 2)) ? (1u << (act & 3)) : 0u);
            if (t + 1 < P) {
                const float e_a = s_e[warp * 128 + act];
                const float v_a = sV[(size_t)act * E + dim0 + kd];
                sub_product(Ohi, Olo, e_a, v_a);
                float zs, ze;
                two_sum(Zhi, -e_a, zs, ze);
                Zhi = zs;
                Zlo = __fadd_rn(Zlo, ze);
                __syncwarp();                           // every lane has read e_a
                if (lane == 0) s_e[warp * 128 + act] = 0.f;     // a repeated (infeasible) action subtracts nothing
                // what is left is a sliver of what the exponentials were scaled for (or cancelled altogether): rescale
                if (!(Zhi > zref * 1.52587890625e-05f)) recompute();
                else if ((g & 1) == 0) so[dim0 + kd] = __fadd_rn(Ohi, Olo);
            }
            SLAPK_TRACE_AT(5);
            ++t;
        }

        // ================================================== epilogue =====================
        SLAPK_TRACE_INST(4);
        __syncthreads();                            // decode done: the value table and the logit keys are free
        if (b + gridDim.x < a.B) {                  // next instance: values -> shared memory, logit keys -> registers
            if (tid == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                stage_values(b + gridDim.x);
            }
            load_logit_keys(b + gridDim.x);
        }
        const int nact = t;
        // chosen log-probs: logp = (x_a - m) - log(sum exp(x - m))   (utils/decoding.py:191,364-368)
        for (int qb = warp; qb < nact; qb += 16 * kWarps) {
            float es[16], mm[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                const int q = qb + u * kWarps;
                es[u] = 0.f;
                mm[u] = 0.f;
                if (q < nact) {
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
            if (lane < 16 && q < nact) {
                const float lpa = (sxs[q * 128 + s_act[q]] - my_m) - logf(my_es);
                s_lp[q] = lpa;
                if (!(lpa > -1000.f)) status |= RL4CO_ST_NEG_INF_LOGP;
            }
        }
        SLAPK_TRACE_INST(5);
        // slap/env.py:131-143: closed tour through the locations of every order (assignment[t] = s_act[t])
        float part = 0.f;
        for (int o = tid; o < a.O; o += kThreads) {
            const int *ord = s_pick + o * a.K;
            const float2 first = s_locs[s_act[ord[0]]];
            float2 prev = first;
            float len = 0.f;
            for (int k = 1; k < a.K; ++k) {
                const float2 cur = s_locs[s_act[ord[k]]];
                len += leg(prev.x, prev.y, cur.x, cur.y);
                prev = cur;
            }
            part += len + leg(prev.x, prev.y, first.x, first.y);
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
        for (int q = tid; q < T; q += kThreads) {    // pad with 0 / log-prob 0 beyond the P assignments
            a.actions[(size_t)r * T + q] = q < nact ? (int64_t)s_act[q] : 0;
            if (a.logp) a.logp[(size_t)r * T + q] = q < nact ? s_lp[q] : 0.f;
        }
        if (a.assignment)
            for (int q = tid; q < a.P; q += kThreads) a.assignment[(size_t)r * a.P + q] = q < nact ? s_act[q] : -1;
        SLAPK_TRACE_INST(6);
        __syncthreads();                         // the next instance rewrites the small arrays
    }
}

__host__ inline int nodes_per_thread(int N) { return N <= 32 ? 2 : N <= 64 ? 4 : N <= 112 ? 7 : 8; }

template <int M>
static int launch(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    const Layout L = make_layout(a.N, a.P < a.T ? a.P : a.T, M, a.O * a.K);
    cudaError_t e = cudaFuncSetAttribute(am_rollout_slap_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
    if (e != cudaSuccess) {
        set_error("am_rollout(slap): cannot reserve %zu B of shared memory: %s", L.total, cudaGetErrorString(e));
        return 2;
    }
    const int64_t cap = (int64_t)2 * sm_count;               // persistent: two CTAs per SM
    const int64_t grid = a.B < cap ? a.B : cap;
    am_rollout_slap_kernel<M><<<(unsigned)grid, kThreads, L.total, stream>>>(a);
    RL4CO_CHECK_LAUNCH("am_rollout(slap)");
    return 0;
}

}  // namespace slapk

bool rollout_slap_supports(const rl4co_am_rollout_args &a) {
    if (a.env != RL4CO_ENV_SLAP || a.multistart || a.S > 1) return false;
#ifndef SLAPK_TRACE
    if (a.logp_all != nullptr) return false;
#endif
    if (a.N > 128) return false;
    const int P = a.P < a.T ? a.P : a.T;
    return slapk::make_layout(a.N, P, slapk::nodes_per_thread(a.N), a.O * a.K).total <= slapk::kSmemLimit;
}

int launch_rollout_slap(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    switch (slapk::nodes_per_thread(a.N)) {
        case 2: return slapk::launch<2>(a, stream);
        case 4: return slapk::launch<4>(a, stream);
        case 7: return slapk::launch<7>(a, stream);
        default: return slapk::launch<8>(a, stream);
    }
}

}  // namespace rl4co
