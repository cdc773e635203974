// K2 (SLAP specialisation) - persistent Attention-Model rollout for the storage-location-assignment env.
//
// What SLAP allows that the routing envs do not: its context embedding returns zeros (`SLAPContext`,
// examples/slap.py:11-84 - the wiring BASELINE.json names), so the decoder query is the projected graph context
// alone (am/decoder.py:141-142) and the glimpse QUERY IS THE SAME AT EVERY STEP of a trajectory, and so are the
// glimpse scores s[h][n] = <q_h, K[n]_h>.  Only the mask changes, and it only ever loses the one location that
// was just assigned (slap/env.py:38-93).  The per-step glimpse (attention.py:289-306)
//     o_h = sum_{n alive} 2^(s[n] - m) V[n]_h / sum_{n alive} 2^(s[n] - m)
// is therefore maintained INCREMENTALLY.  The numerator is defined as the sum of the individually rounded products
// fl(e_n V[n]) (what an fp32 implementation adds up anyway) and, like the denominator, is held as a two-float value
// hi + lo: built with error-free TwoSum additions, and reduced at every step by adding -fl(e_a V[a]), -e_a of the
// assigned location the same way - which removes exactly what the build put in.  What remains is therefore the sum
// over the alive locations to ~2^-44 of the sum it STARTED from, i.e. still to fp32 accuracy after the remainder
// has shrunk by many orders of magnitude (a plain fp32 starting sum would not do: its rounding error is small
// against the full sum only, and the first dominant location to leave would expose it).  When what is left of a
// head's denominator falls below 2^-16 of its value at the last scaling (the head was concentrated on locations that
// are gone; exponentials of the others may have been flushed to zero), every head is rebuilt from the stored scores
// with the maximum of the nodes still alive - the softmax shift the reference uses at that step.  Any shift gives the
// same quotient; the exponentials differ from the step-by-step kernel's by the rounding of ex2.approx (2^-22), the
// same class of error that kernel has.
//
// Per step that leaves ONE mat-vec, the pointer logits <o, L'[n]> (project_out folded into L', attention.py:
// 276-282), and a step is a latency chain, not arithmetic.  Ownership is therefore NODE-major, so that a step has
// a single block barrier and no exchange of partial logits through shared memory:
//   warp w owns the NW consecutive nodes w NW .. w NW + NW - 1; lane l owns dims 4l .. 4l+3 (head l >> 2) of their
//   logit keys, in registers for the whole trajectory (NW = 13 at N = 100: 52 registers);
//   every warp keeps the whole glimpse state redundantly, lane l its own 4 dims of O and its head's Z;
//   logits: NW register dot products per lane, one 32-lane transposing butterfly (16 shuffles) that leaves the
//   full logit of node w NW + (l >> 1) in lane l; tanh clip, mask, (Gumbel noise), warp arg-max by two redux;
//   the 8 warp winners meet in shared memory (double buffered), ONE barrier, every warp picks the block winner;
//   update of O, Z from shared memory: e[h][a] (formed once per scaling) and the value row V[a] (the instance's
//   value table is landed in shared memory by TMA one instance ahead; logit keys are loaded one instance ahead
//   into the registers the finished instance no longer needs).
// Two CTAs per SM (about 84 KB of shared memory and 128 registers each) cover each other's chains.
//
// Selection, Gumbel noise keying, log-prob bookkeeping and outputs follow am_rollout_rf.cu, so for the same seed a
// sampled rollout picks the same locations unless a logit differs in the last bits.  Restricted to the standard
// logit post-processing (tanh clipping, temperature 1, masked logits); anything else runs on am_rollout_rf.cu.
//
// Reference op order being reproduced: SURVEY.md appendix A; rl4co/models/zoo/am/decoder.py:134-200,
// rl4co/models/nn/attention.py:263-309, rl4co/utils/decoding.py:141-191,327-397,
// rl4co/envs/warehousing/slap/env.py:38-143.
#include <stdlib.h>

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
constexpr size_t kSmemLimit = 112 * 1024;   // two CTAs per SM

// Shared-memory layout with COMPILE-TIME offsets (sized for the limits below): with run-time offsets the decode loop spent
// ~50 of its 284 instructions per step rebuilding shared-memory addresses the 128-register cap would not let it keep
constexpr int kMaxSteps = 32;        // P: products = decode steps
constexpr int kMaxPick = 256;        // O * K: picklist entries
struct Layout {
    static constexpr size_t off_xs = 0;                                          // processed logits, one row per step
    static constexpr size_t off_sc = off_xs + (size_t)kMaxSteps * 128 * 4;      // glimpse scores [head][node]
    static constexpr size_t off_e = off_sc + (size_t)kWarps * 128 * 4;          // e[head][node] = 2^(s - m) at the last scaling
    static constexpr size_t off_po = off_e + (size_t)kWarps * 128 * 4;          // per-warp partial numerators, hi and lo parts
    static constexpr size_t off_z = off_po + (size_t)2 * kWarps * E * 4;        // glimpse denominators [head], (hi, lo)
    static constexpr size_t off_best = off_z + 16 * 4;                          // warp winners (value, node), double buffered
    static constexpr size_t off_mw = off_best + 2 * kWarps * 8;
    static constexpr size_t off_act = off_mw + kMaxSteps * 4;
    static constexpr size_t off_ain = off_act + kMaxSteps * 4;
    static constexpr size_t off_lp = off_ain + kMaxSteps * 4;
    static constexpr size_t off_locs = off_lp + kMaxSteps * 4;
    static constexpr size_t off_noise = off_locs + 128 * 8;                     // sampling: Gumbel noise of 16 steps [16][node]
    static constexpr size_t off_part = off_noise + 16 * 128 * 4;
    static constexpr size_t off_pick = off_part + 64 * 4;                       // picklist as int
    static constexpr size_t off_bar = off_pick + kMaxPick * 4;
    static constexpr size_t off_v = (off_bar + 16 + 127) / 128 * 128;           // glimpse values [node][128], landed by TMA
    static constexpr size_t total = off_v + (size_t)128 * E * 4;
};
static_assert(Layout::total <= kSmemLimit, "two CTAs per SM");

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
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) {
    u64 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
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
// a lane's 4 dims of one table row: 16 bytes, a warp reads the whole 512 B row in one request
__device__ __forceinline__ void ldg_128(const float *p, u64 (&v)[2]) {
    asm volatile("ld.global.nc.v2.u64 {%0, %1}, [%2];" : "=l"(v[0]), "=l"(v[1]) : "l"(p));
}
__device__ __forceinline__ void prefetch_l2(const void *p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
// error-free pieces of the compensated (two-float) accumulators: value = hi + lo
__device__ __forceinline__ void two_sum(float a, float b, float &s, float &err) {
    s = __fadd_rn(a, b);
    const float bb = __fsub_rn(s, a);
    err = __fadd_rn(__fsub_rn(a, __fsub_rn(s, bb)), __fsub_rn(b, bb));
}
__device__ __forceinline__ u64 fadd2(u64 a, u64 b) {
    u64 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
// (hi, lo) += q on two dims at once (packed fp32 pairs, m1 = (-1, -1)): TwoSum, with a - b written fma(b, -1, a), which is
// exact.  The glimpse numerator is DEFINED as the sum of the individually rounded products fl(e_n V_n): a location is
// taken out by adding -fl(e_a V_a), which removes exactly what the build put in
__device__ __forceinline__ void add_term2(u64 &hi, u64 &lo, u64 q, u64 m1) {
    const u64 s = fadd2(hi, q);
    const u64 bb = ffma2(hi, m1, s);                // s - hi
    const u64 t1 = ffma2(bb, m1, s);                // s - bb
    const u64 t2 = ffma2(t1, m1, hi);               // hi - (s - bb)
    const u64 t3 = ffma2(bb, m1, q);                // q - bb
    hi = s;
    lo = fadd2(lo, fadd2(t2, t3));
}

// -DSLAPK_TRACE (tools/trace_slap.py): thread 0 of CTA 0 stamps clock64() into a.logp_all, reused as a [P + 1][8] int64
// buffer, during its SECOND instance (the other CTA of the SM is then in flight): five points per step, and in row P the
// phases of the instance
#ifdef SLAPK_TRACE
#define SLAPK_TRACE_AT(k) do { if (tid == 0 && blockIdx.x == 0 && it == 1) reinterpret_cast<long long *>(a.logp_all)[t * 8 + (k)] = clock64(); } while (0)
#define SLAPK_TRACE_INST(k) do { if (tid == 0 && blockIdx.x == 0 && it == 1) reinterpret_cast<long long *>(a.logp_all)[P * 8 + (k)] = clock64(); } while (0)
#define SLAPK_TRACE_BUILD(k) do { if (tid == 0 && blockIdx.x == 0 && it == 1 && t == 0) reinterpret_cast<long long *>(a.logp_all)[(P + 1) * 8 + (k)] = clock64(); } while (0)
#else
#define SLAPK_TRACE_AT(k) do { } while (0)
#define SLAPK_TRACE_INST(k) do { } while (0)
#define SLAPK_TRACE_BUILD(k) do { } while (0)
#endif

constexpr int pow2_at_least(int v) { return v <= 4 ? 4 : v <= 8 ? 8 : 16; }
constexpr int log2_of(int v) { return v == 4 ? 2 : v == 8 ? 3 : 4; }

template <int NW>
__global__ void __launch_bounds__(kThreads, 2)
am_rollout_slap_kernel(const rl4co_am_rollout_args a) {
    constexpr int PADV = pow2_at_least(NW);         // butterfly width
    constexpr int SH = 5 - log2_of(PADV);           // lane l ends up with the logit of the warp's node l >> SH
    extern __shared__ __align__(128) unsigned char smem[];
    const int N = a.N, T = a.T, P = a.P < a.T ? a.P : a.T;
    const int n_pick = a.O * a.K;
    using L = Layout;
    float *sxs = reinterpret_cast<float *>(smem + L::off_xs);
    float *s_sc = reinterpret_cast<float *>(smem + L::off_sc);
    float *s_e = reinterpret_cast<float *>(smem + L::off_e);
    float *s_po = reinterpret_cast<float *>(smem + L::off_po);
    float *s_z = reinterpret_cast<float *>(smem + L::off_z);
    float2 *s_best = reinterpret_cast<float2 *>(smem + L::off_best);
    float *s_mw = reinterpret_cast<float *>(smem + L::off_mw);
    int *s_act = reinterpret_cast<int *>(smem + L::off_act);
    int *s_ain = reinterpret_cast<int *>(smem + L::off_ain);
    float *s_lp = reinterpret_cast<float *>(smem + L::off_lp);
    float2 *s_locs = reinterpret_cast<float2 *>(smem + L::off_locs);
    float4 *s_noise = reinterpret_cast<float4 *>(smem + L::off_noise);
    float *spart = reinterpret_cast<float *>(smem + L::off_part);
    int *s_pick = reinterpret_cast<int *>(smem + L::off_pick);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + L::off_bar);
    float *sV = reinterpret_cast<float *>(smem + L::off_v);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int d0 = 4 * lane;                        // first of the 4 dims this lane owns
    const int hl = lane >> 2;                       // their head
    const int n0 = warp * NW;                       // first node of this warp
    const int vi = lane >> SH;                      // the warp's node whose logit the butterfly leaves in this lane
    const int n_me = n0 + vi;
    const bool valid_me = vi < NW && n_me < N;
    const bool writer_me = valid_me && (lane & ((1 << SH) - 1)) == 0;
    const int64_t Mrows = a.B * (int64_t)N;         // rows per table
    const size_t tab_bytes = (size_t)N * E * 4;
    const float clip = a.tanh_clipping;
    const bool sampling = a.mode == RL4CO_MODE_SAMPLING, greedy = a.mode == RL4CO_MODE_GREEDY;

    unsigned rown = 0;                              // reader layout: bit k of lane l <-> node 4 l + k
#pragma unroll
    for (int k = 0; k < 4; ++k) rown |= (4 * lane + k < N) ? (1u << k) : 0u;

    // pad columns of the logit rows are never written again: -inf once
    for (int q = tid; q < P * 128; q += kThreads)
        if ((q & 127) >= N) sxs[q] = -INFINITY;
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
    u64 L2[NW][2];                                  // logit keys of the current instance, loaded one instance ahead
    auto load_logit_keys = [&](int64_t b) {
        const float *lb = a.tables + ((size_t)2 * Mrows + b * (int64_t)N) * E + d0;
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            if (n0 + i < N) ldg_128(lb + (size_t)(n0 + i) * E, L2[i]);
            else { L2[i][0] = 0ull; L2[i][1] = 0ull; }
        }
    };
    // small per-instance inputs, also fetched one instance ahead: location of node tid, picklist entry tid, this lane's 4 dims
    // of the graph context
    float2 nx_loc = make_float2(0.f, 0.f);
    long long nx_pick = 0;
    float4 nx_g = make_float4(0.f, 0.f, 0.f, 0.f);
    auto load_small = [&](int64_t b) {
        if (tid < N) nx_loc = __ldg(reinterpret_cast<const float2 *>(a.locs + b * (int64_t)N * 2) + tid);
        if (tid < n_pick) nx_pick = __ldg(a.picklist + b * (int64_t)n_pick + tid);
        if (a.gctx) nx_g = __ldg(reinterpret_cast<const float4 *>(a.gctx + b * E + d0));
    };
    if ((int64_t)blockIdx.x < a.B) {
        if (tid == 0) stage_values(blockIdx.x);
        load_logit_keys(blockIdx.x);
        load_small(blockIdx.x);
    }

    int it = 0;
    for (int64_t b = blockIdx.x; b < a.B; b += gridDim.x, ++it) {
        const int64_t r = b;
        int t = 0, status = 0;
#ifdef SLAPK_TRACE
        int n_rebuilds = 0;
#endif
        SLAPK_TRACE_INST(0);
        // ------------------------------------------------ glimpse scores, once (attention.py:289-306), base-2 scaled
        {
            // the one query of the trajectory, pre-scaled (zero without a graph context)
            const u64 q2[2] = {pack2(nx_g.x * kScoreScale, nx_g.y * kScoreScale), pack2(nx_g.z * kScoreScale, nx_g.w * kScoreScale)};
            const float *kb = a.tables + ((size_t)0 * Mrows + b * (int64_t)N) * E + d0;
            u64 kr[NW][2];
#pragma unroll
            for (int i = 0; i < NW; ++i) {
                if (n0 + i < N) ldg_128(kb + (size_t)(n0 + i) * E, kr[i]);
                else { kr[i][0] = 0ull; kr[i][1] = 0ull; }
            }
            float sv[NW];
#pragma unroll
            for (int i = 0; i < NW; ++i) sv[i] = hsum2(ffma2(kr[i][1], q2[1], fmul2(kr[i][0], q2[0])));
#pragma unroll
            for (int i = 0; i < NW; ++i) sv[i] += __shfl_xor_sync(kFull, sv[i], 1);
#pragma unroll
            for (int i = 0; i < NW; ++i) sv[i] += __shfl_xor_sync(kFull, sv[i], 2);
            if ((lane & 3) == 0) {
#pragma unroll
                for (int i = 0; i < NW; ++i)
                    if (n0 + i < N) s_sc[hl * 128 + n0 + i] = sv[i];
            }
        }
        SLAPK_TRACE_INST(1);
        // small per-instance state -> shared memory (the previous instance's epilogue ended with a barrier)
        {
            if (tid < N) s_locs[tid] = nx_loc;
            if (tid < n_pick) s_pick[tid] = (int)nx_pick;       // n_pick <= kMaxPick = kThreads
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
        unsigned ralive = rown;
        if (lane == 0) ralive &= ~1u;
        bool alive_me = valid_me && n_me != 0;

        // ------------------------------------------------ glimpse state, replicated in every warp: this lane's 4 dims of
        // the numerator and its head's denominator as two-float sums; zref = denominator at the last scaling
        u64 Ohi[2], Olo[2];                         // dims (4l, 4l+1) and (4l+2, 4l+3), packed
        float Zhi, Zlo, zref;
        // (re)build it from the scores with the maximum of the alive nodes as the shift.  Block-wide: warp w = head w forms
        // the exponentials and the denominator, then every warp sums e V over its own nodes, then all partials are added
        // Everything is summed as two-float values: the state is later REDUCED by exact subtractions, so an error that is
        // small against the full sum would not be small against what remains once the dominant locations are gone
        auto build_state = [&]() {
            const u64 m1 = pack2(-1.f, -1.f);
            {
                const float4 s4 = *reinterpret_cast<const float4 *>(s_sc + warp * 128 + 4 * lane);
                const float s0 = (ralive & 1u) ? s4.x : -INFINITY, s1 = (ralive & 2u) ? s4.y : -INFINITY;
                const float s2 = (ralive & 4u) ? s4.z : -INFINITY, s3 = (ralive & 8u) ? s4.w : -INFINITY;
                const float mx = redux_max(fmaxf(fmaxf(s0, s1), fmaxf(s2, s3)));
                const float4 e4 = make_float4(ex2f(s0 - mx), ex2f(s1 - mx), ex2f(s2 - mx), ex2f(s3 - mx));   // 2^-inf = 0
                *reinterpret_cast<float4 *>(s_e + warp * 128 + 4 * lane) = e4;
                float zh, zl, err;
                two_sum(e4.x, e4.y, zh, zl);
                two_sum(zh, e4.z, zh, err); zl += err;
                two_sum(zh, e4.w, zh, err); zl += err;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {      // symmetric, so every lane ends with the same pair
                    const float oh = __shfl_xor_sync(kFull, zh, o), ol = __shfl_xor_sync(kFull, zl, o);
                    two_sum(zh, oh, zh, err);
                    zl = (zl + ol) + err;
                }
                if (lane == 0) reinterpret_cast<float2 *>(s_z)[warp] = make_float2(zh, zl);
            }
            SLAPK_TRACE_BUILD(1);
            __syncthreads();
            SLAPK_TRACE_BUILD(2);
            {
                u64 ph[2] = {0ull, 0ull}, pl[2] = {0ull, 0ull};
#pragma unroll
                for (int i = 0; i < NW; ++i) {
                    const int n = n0 + i < N ? n0 + i : N - 1;      // absent nodes: e = 0, any row
                    const float e = s_e[hl * 128 + n0 + i];
                    const ulonglong2 v4 = *reinterpret_cast<const ulonglong2 *>(sV + (size_t)n * E + d0);
                    const u64 e2 = pack2(e, e);
                    add_term2(ph[0], pl[0], fmul2(e2, v4.x), m1);
                    add_term2(ph[1], pl[1], fmul2(e2, v4.y), m1);
                }
                *reinterpret_cast<ulonglong2 *>(s_po + warp * E + d0) = make_ulonglong2(ph[0], ph[1]);
                *reinterpret_cast<ulonglong2 *>(s_po + (kWarps + warp) * E + d0) = make_ulonglong2(pl[0], pl[1]);
            }
            SLAPK_TRACE_BUILD(3);
            __syncthreads();
            SLAPK_TRACE_BUILD(4);
            Ohi[0] = Ohi[1] = Olo[0] = Olo[1] = 0ull;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {      // (splitting these chains in two did not help: the build competes for
                const ulonglong2 h4 = *reinterpret_cast<const ulonglong2 *>(s_po + w * E + d0);     // issue slots, not latency)
                const ulonglong2 l4 = *reinterpret_cast<const ulonglong2 *>(s_po + (kWarps + w) * E + d0);
                add_term2(Ohi[0], Olo[0], h4.x, m1);
                add_term2(Ohi[1], Olo[1], h4.y, m1);
                Olo[0] = fadd2(Olo[0], l4.x);
                Olo[1] = fadd2(Olo[1], l4.y);
            }
            const float2 z2 = reinterpret_cast<const float2 *>(s_z)[hl];
            Zhi = z2.x; Zlo = z2.y; zref = Zhi;
            SLAPK_TRACE_BUILD(5);
        };
        SLAPK_TRACE_INST(2);
        mbar_wait(bar, par);                        // value table landed in shared memory
        par ^= 1u;
        SLAPK_TRACE_BUILD(6);
        __syncthreads();                            // scores / noise / given actions visible to every warp
        SLAPK_TRACE_BUILD(0);
        build_state();
        SLAPK_TRACE_INST(3);

        // ================================================== decode loop ==================
        while (t < P) {
            SLAPK_TRACE_AT(0);
            float Gn = 0.f;
            if (sampling) {                             // noise ring as in am_rollout_rf.cu, read by node
                if ((t & 7) == 0 && t > 0) gen_noise(t + 8 + warp);
                Gn = reinterpret_cast<const float *>(s_noise)[(t & 15) * 128 + (valid_me ? n_me : 0)];
            }
            // ---- pointer logits of the own nodes, project_out folded (attention.py:276-282): o = O / Z, then <o, L'[n]>
            float tot;
            {
                const float inv = rcpf(Zhi + Zlo);          // softmax normaliser of this lane's head
                const u64 inv2 = pack2(inv, inv);
                const u64 o2[2] = {fmul2(fadd2(Ohi[0], Olo[0]), inv2), fmul2(fadd2(Ohi[1], Olo[1]), inv2)};
                float v[PADV];
#pragma unroll
                for (int i = 0; i < PADV; ++i) v[i] = i < NW ? hsum2(ffma2(L2[i < NW ? i : 0][1], o2[1], fmul2(L2[i < NW ? i : 0][0], o2[0]))) : 0.f;
                // transposing butterfly: after the stage with offset o a lane keeps the half of the values selected by
                // its bit o; the last stages are plain exchanges.  Lane l ends with the total of value l >> SH
                int half = PADV / 2;
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) {
                    if (half >= 1) {
                        const bool up = (lane & o) != 0;
#pragma unroll
                        for (int k = 0; k < PADV / 2; ++k) {
                            if (k < half) {
                                const float send = up ? v[k] : v[k + half], keep = up ? v[k + half] : v[k];
                                v[k] = keep + __shfl_xor_sync(kFull, send, o);
                            }
                        }
                        half >>= 1;
                    } else {
                        v[0] += __shfl_xor_sync(kFull, v[0], o);
                    }
                }
                tot = v[0];
            }
            SLAPK_TRACE_AT(1);
            // ---- process_logits (utils/decoding.py:141-191): clip * tanh(z) = clip - 2 clip / (2^(2 log2(e) z) + 1), mask
            float xv;
            {
                const float e2 = ex2f(tot * (kInvSqrtE * 2.f * kLog2e));
                xv = fmaf(-2.f * clip, rcpf(e2 + 1.f), clip);
                if (valid_me && tot != tot) status |= RL4CO_ST_NAN_LOGITS;
                xv = alive_me ? xv : -INFINITY;
            }
            if (writer_me) sxs[t * 128 + n_me] = xv;    // the step's log-prob is formed after the loop
            // ---- selection (utils/decoding.py:371-397,410-499)
            int act;
            float gm = clip;                            // log-sum-exp shift: the bound of the clipped logits, or their maximum
            if (a.mode == RL4CO_MODE_EVALUATE) {
                act = s_ain[t];
                if (act < 0) { status |= RL4CO_ST_ACTION_RANGE; act = 0; }
            } else {
                // greedy: arg-max of the logits; sampling: Gumbel-max, arg-max of logits + noise (rollout_util.cuh); lowest
                // node on ties.  Warp winner by two redux, block winner after the step's only barrier
                const float y = valid_me ? xv + Gn : -INFINITY;
                const float ym = redux_max(y);
                const unsigned win = redux_min_u32((valid_me && y == ym) ? (unsigned)n_me : 0x7fffffffu);
                float2 *slot = s_best + (t & 1) * kWarps;
                if (lane == 0) slot[warp] = make_float2(ym, __uint_as_float(win));
                SLAPK_TRACE_AT(2);
                __syncthreads();
                SLAPK_TRACE_AT(3);
                // lane w < 8 takes warp w's winner: the largest value, the lowest node among equals
                const float2 mine = slot[lane & 7];
                const float best = redux_max(mine.x);
                const unsigned bi = redux_min_u32(mine.x == best ? __float_as_uint(mine.y) : 0x7fffffffu);
                act = bi < (unsigned)N ? (int)bi : 0;
                if (greedy) gm = best;
            }
            // feasibility of the chosen action (utils/decoding.py:377-379,393-395): the lane that holds its logit checks
            if ((n_me == act && vi < NW && !alive_me) || act >= N) status |= RL4CO_ST_INFEASIBLE_ACTION;
            if (tid == 0) { s_act[t] = act; s_mw[t] = gm; }
            // ---- transition (slap/env.py:38-93): the location leaves the mask and every head's glimpse state
            alive_me = alive_me && n_me != act;
            ralive &= ~((lane == (act >> 2)) ? (1u << (act & 3)) : 0u);
            if (t + 1 < P) {
                const float e_a = s_e[hl * 128 + act];
                const ulonglong2 va = *reinterpret_cast<const ulonglong2 *>(sV + (size_t)act * E + d0);
                const u64 ne2 = pack2(-e_a, -e_a), m1 = pack2(-1.f, -1.f);
                add_term2(Ohi[0], Olo[0], fmul2(ne2, va.x), m1);    // -fl(e_a V_a): the term the build added, negated
                add_term2(Ohi[1], Olo[1], fmul2(ne2, va.y), m1);
                float zs, ze;
                two_sum(Zhi, -e_a, zs, ze);
                Zhi = zs;
                Zlo = __fadd_rn(Zlo, ze);
                // what is left of some head is a sliver (2^-16) of what its exponentials were scaled for, or cancelled
                // altogether: rebuild with the maximum of what is alive.  Every warp holds the same state, so all of them take the branch together
                if (__any_sync(kFull, !(Zhi > zref * 1.52587890625e-05f))) {
                    __syncthreads();                    // everyone is done with e[] and the partials of the last build
                    build_state();
#ifdef SLAPK_TRACE
                    ++n_rebuilds;
#endif
                }
            }
            SLAPK_TRACE_AT(4);
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
            load_small(b + gridDim.x);
        }
        const int nact = t;
        // chosen log-probs: logp = (x_a - m) - log(sum exp(x - m))   (utils/decoding.py:191,364-368): 8 threads per step, 16 nodes
        // each, 32 steps per pass
        for (int q0 = 0; q0 < nact; q0 += kThreads / 8) {
            const int q = q0 + (tid >> 3), part = tid & 7;
            float es = 0.f, m = 0.f;
            if (q < nact) {
                m = s_mw[q];
                const float4 *row = reinterpret_cast<const float4 *>(sxs + q * 128 + 16 * part);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float4 x4 = row[u];
                    es += (ex2f((x4.x - m) * kLog2e) + ex2f((x4.y - m) * kLog2e)) +
                          (ex2f((x4.z - m) * kLog2e) + ex2f((x4.w - m) * kLog2e));     // masked logits: 2^-inf = 0
                }
            }
            es += __shfl_xor_sync(kFull, es, 1);
            es += __shfl_xor_sync(kFull, es, 2);
            es += __shfl_xor_sync(kFull, es, 4);
            if (q < nact && part == 0) {
                const float lpa = (sxs[q * 128 + s_act[q]] - m) - logf(es);
                s_lp[q] = lpa;
                if (!(lpa > -1000.f)) status |= RL4CO_ST_NEG_INF_LOGP;
            }
        }
        SLAPK_TRACE_INST(5);
        // slap/env.py:131-143: closed tour through the locations of every order (assignment[t] = s_act[t])
        float part = 0.f;
        for (int idx = tid; idx < n_pick; idx += kThreads) {    // one leg per thread: entry k of its order -> entry k + 1 (cyclic)
            const int k = idx % a.K, nxt = k + 1 == a.K ? idx - k : idx + 1;
            const float2 p = s_locs[s_act[s_pick[idx]]], w = s_locs[s_act[s_pick[nxt]]];
            part += leg(p.x, p.y, w.x, w.y);
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
#ifdef SLAPK_TRACE
                if (a.steps) a.steps[r] = nact + (n_rebuilds << 8);     // tools/trace_slap.py: rebuilds per instance
#else
                if (a.steps) a.steps[r] = nact;
#endif
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
        // no barrier here: what the next instance writes before its own barriers (scores, locations, picklist, noise, given
        // actions) is not read by the output loops above, and the reward loop that reads locations / picklist ended before
        // the barrier in the middle of this epilogue
    }
}

__host__ inline int nodes_per_warp(int N) { return N <= 32 ? 4 : N <= 64 ? 8 : N <= 104 ? 13 : 16; }

template <int NW>
static int launch(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    using L = Layout;
    cudaError_t e = cudaFuncSetAttribute(am_rollout_slap_kernel<NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::total);
    if (e == cudaSuccess)   // two CTAs of ~84 KB: ask for the largest shared-memory carve-out
        e = cudaFuncSetAttribute(am_rollout_slap_kernel<NW>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) {
        set_error("am_rollout(slap): cannot reserve %zu B of shared memory: %s", L::total, cudaGetErrorString(e));
        return 2;
    }
    if (getenv("RL4CO_DEBUG_OCCUPANCY")) {
        int nb = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, am_rollout_slap_kernel<NW>, kThreads, L::total);
        cudaFuncAttributes fa;
        cudaFuncGetAttributes(&fa, am_rollout_slap_kernel<NW>);
        fprintf(stderr, "am_rollout(slap)<%d>: %d CTAs/SM, %d regs, %zu B dynamic smem, carveout %d\n", NW, nb, fa.numRegs, L::total,
                fa.preferredShmemCarveout);
    }
    int64_t cap = (int64_t)2 * sm_count;                     // persistent: two CTAs per SM
    if (const char *dbg = getenv("RL4CO_DEBUG_SLAP_GRID")) cap = atoll(dbg) > 0 ? atoll(dbg) : cap;   // timing experiments only
    const int64_t grid = a.B < cap ? a.B : cap;
    am_rollout_slap_kernel<NW><<<(unsigned)grid, kThreads, L::total, stream>>>(a);
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
    if (!(a.tanh_clipping > 0.f && a.temperature == 1.f && a.mask_logits != 0)) return false;
    const int P = a.P < a.T ? a.P : a.T;
    return P <= slapk::kMaxSteps && a.O * a.K <= slapk::kMaxPick;     // the limits of the fixed shared-memory layout
}

int launch_rollout_slap(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    switch (slapk::nodes_per_warp(a.N)) {
        case 4: return slapk::launch<4>(a, stream);
        case 8: return slapk::launch<8>(a, stream);
        case 13: return slapk::launch<13>(a, stream);
        default: return slapk::launch<16>(a, stream);
    }
}

}  // namespace rl4co
