// K2 (TSP, greedy or sampling, single start) - the register-file rollout of am_rollout_rf.cu specialised
// for the headline configuration and extended with NODE COMPACTION.
//
// Same ownership and the same per-step arithmetic as am_rollout_rf.cu (warp = attention head,
// lane = (half head, node group), K / V / folded-L rows of the thread's nodes in registers,
// FFMA2 mat-vecs, two block barriers per step).  What is different:
//   * everything the generic kernel decides at run time (environment, decoding mode, logit
//     post-processing variant, whether the per-step logits fit in shared memory) is fixed here:
//     TSP (tsp/env.py:67-120), greedy or sampling as a template parameter (utils/decoding.py:371-397,410-449), tanh clipping > 0,
//     temperature 1, masked logits, logits of every step kept for the log-probabilities of the
//     epilogue.  Anything else is routed to am_rollout_rf.cu / am_rollout.cu by the dispatcher.
//   * a TSP tour never returns to a visited node, so the rows of visited nodes are dead weight in
//     the register file.  The trajectory runs in stages of 7 (or 8), 4 and 2 rows of 16 node slots:
//     when the number of unvisited nodes drops to 64 and to 32, every warp re-packs the K / V / L
//     rows of its head's surviving nodes (through a 4 KB per-warp scratch, no block barrier) into
//     the first slots, in node order - so "lowest slot" is still "lowest node index" for the argmax
//     tie rule - and continues with a loop body instantiated for fewer rows.  Measured on B200
//     (tools/trace_rollout.py): 1943 clk/step with 7 rows, 1472 with 4, 1306 with 2.
//   * slots are translated back to node ids inside the argmax reduction itself: the key of
//     redux.sync.min is (slot << 8 | node), so the winning lane's node id costs no extra exchange.
//
// Reference op order being reproduced (all fp32): SURVEY.md appendix A;
// rl4co/models/zoo/am/decoder.py:134-200, rl4co/models/nn/attention.py:263-309,
// rl4co/utils/decoding.py:141-191,327-397, rl4co/envs/routing/tsp/env.py:67-173.
#include "rollout_util.cuh"

namespace rl4co {
namespace tspg {

typedef unsigned long long u64;

constexpr int E = RL4CO_EMBED_DIM;   // 128
constexpr int kThreads = 256;
constexpr int kWarps = 8;
constexpr int kMaxT = 128;            // steps (= nodes) the fixed part of the layout is sized for
constexpr float kScoreScale = 0.25f * 1.4426950408889634f;   // 1/sqrt(16) * log2(e)
constexpr float kInvSqrtE = 0.08838834764831845f;             // 1/sqrt(128)
constexpr float kLog2e = 1.4426950408889634f;
constexpr int PLS = 144;              // row stride of the partial-logit matrix [head][slot], = 16 (mod 32)
#ifndef RL4CO_TSPG_FUSED_ROWS
#define RL4CO_TSPG_FUSED_ROWS 2   // measured: fusing the 4-row stage (2 slots per lane) is 1240 clk/step against 1230
#endif
constexpr int kScrSlots = 64;         // compaction target: at most 4 rows of 16 slots

// Shared memory.  The buffers touched on every step sit at compile-time offsets.
constexpr size_t OFF_PL = 0;                                        // partial logits [2][head][slot] (2nd copy: fused stage)
constexpr size_t OFF_O = OFF_PL + (size_t)2 * kWarps * PLS * 4;     // glimpse heads (warp private)
constexpr size_t OFF_ZERO = OFF_O + E * 4;                          // 512 B of zeros: the "context row" of step 0
constexpr size_t OFF_MW = OFF_ZERO + E * 4;                         // per step: max logit
constexpr size_t OFF_ACT = OFF_MW + kMaxT * 4;                      // per step: chosen node
constexpr size_t OFF_SLOT = OFF_ACT + kMaxT * 4;                    // per step: its slot (index into the kept logits)
constexpr size_t OFF_LP = OFF_SLOT + kMaxT * 4;                     // per step: chosen log-prob
constexpr size_t OFF_NODE = OFF_LP + kMaxT * 4;                     // per warp: slot -> node id [8][128]
constexpr size_t OFF_LOCS = OFF_NODE + (size_t)kWarps * 128 * 4;    // coordinates [128] float2
constexpr size_t OFF_PART = OFF_LOCS + 128 * 8;
constexpr size_t OFF_BAR = OFF_PART + 64 * 4;
constexpr size_t OFF_SCR = OFF_BAR + 16;                            // per warp compaction scratch [64 slots][16 dims]
constexpr size_t OFF_NOISE = OFF_SCR + (size_t)kWarps * kScrSlots * 16 * 4;  // sampling: Gumbel noise of 16 steps [16][128]
constexpr size_t OFF_XS = OFF_NOISE + (size_t)16 * 128 * 4;        // kept logits [T][128]
__host__ __device__ inline size_t off_c(int T) { return OFF_XS + (size_t)T * 128 * 4; }      // context table x2
__host__ __device__ inline size_t smem_total(int N, int T) { return off_c(T) + (size_t)2 * N * E * 4; }

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
// 32 contiguous bytes in one request (SASS: LDG.E.256): a thread's 8 dims of one table row are exactly one sector
__device__ __forceinline__ void ldg_256(const float *p, u64 (&v)[4]) {
    asm volatile("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(v[0]), "=l"(v[1]), "=l"(v[2]), "=l"(v[3]) : "l"(p));
}
__device__ __forceinline__ void lds_2x128(uint32_t addr, u64 (&v)[4]) {
    asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(v[0]), "=l"(v[1]) : "r"(addr));
    asm volatile("ld.shared.v2.u64 {%0, %1}, [%2+16];" : "=l"(v[2]), "=l"(v[3]) : "r"(addr));
}
__device__ __forceinline__ void sts_2x128(uint32_t addr, const u64 (&v)[4]) {
    asm volatile("st.shared.v2.u64 [%0], {%1, %2};" ::"r"(addr), "l"(v[0]), "l"(v[1]) : "memory");
    asm volatile("st.shared.v2.u64 [%0+16], {%1, %2};" ::"r"(addr), "l"(v[2]), "l"(v[3]) : "memory");
}

// -DRF_TRACE: thread 0 of CTA 0 stamps clock64() at eight points of every step of its first instance
#ifdef RF_TRACE
#define TG_TRACE_AT(k) do { if (threadIdx.x == 0 && blockIdx.x == 0 && c.trace) reinterpret_cast<long long *>(c.trace)[c.t * 8 + (k)] = clock64(); } while (0)
#define TG_TRACE_ROW(row, k) do { if (threadIdx.x == 0 && blockIdx.x == 0 && c.trace) reinterpret_cast<long long *>(c.trace)[(row) * 8 + (k)] = clock64(); } while (0)
#else
#define TG_TRACE_AT(k) do { } while (0)
#define TG_TRACE_ROW(row, k) do { } while (0)
#endif

// per-thread state carried through the stages of one trajectory
struct Ctx {
    uint32_t c_addr;        // shared address of this thread's 8 dims of context row 0 (current buffer)
    uint32_t x_addr;        // shared address of the kept-logit row of step t
    u64 qbs2[4];            // constant part of the query, pre-scaled
    unsigned alive, own;    // owner layout: bit j <-> slot g + 16 j
    unsigned candk[4];      // (slot << 8 | node) of the reader-layout slots 4 lane .. 4 lane + 3
    int cur, t, status;
    float clip;
    uint64_t seed, offset, r;   // sampling: Philox key, counter base, rollout index (utils/decoding.py:381-397)
    void *trace;
    int trow;               // RF_TRACE: first spare row of the trace buffer
};

// Sampling noise.  Every warp needs the same 128 Gumbel values per step; instead of eight redundant Philox calls
// per step, warp w draws the noise of step t0 + 8 + w once every eight steps (t0 = multiple of 8) into a ring of
// 16 steps in shared memory.  A row is rewritten only after the barrier(s) of the step that read it.
__device__ __forceinline__ void gen_noise(const Ctx &c, int step, unsigned char *smem) {
    float Gn[4];
    gumbel4(c.seed, c.offset, c.r, (uint32_t)step, threadIdx.x & 31, Gn);
    reinterpret_cast<float4 *>(smem + OFF_NOISE)[(step & 15) * 32 + (threadIdx.x & 31)] = make_float4(Gn[0], Gn[1], Gn[2], Gn[3]);
}

// ---------------------------------------------------------------------------------------------
// decode steps t .. t_stop-1 with MM rows of 16 slots
// ---------------------------------------------------------------------------------------------
template <int MM, bool SAMPLING>
__device__ __forceinline__ void run_stage(Ctx &c, const u64 (&K2)[MM][4], const u64 (&V2)[MM][4], const u64 (&L2)[MM][4],
                                          const int t_stop, unsigned char *smem) {
    float *spl = reinterpret_cast<float *>(smem + OFF_PL);
    float *so = reinterpret_cast<float *>(smem + OFF_O);
    float *s_mw = reinterpret_cast<float *>(smem + OFF_MW);
    int *s_act = reinterpret_cast<int *>(smem + OFF_ACT);
    int *s_slot = reinterpret_cast<int *>(smem + OFF_SLOT);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int hh = lane >> 4, g = lane & 15;
    const int dim0 = 16 * warp + 8 * hh;
    const u64 scale2 = pack2(kScoreScale, kScoreScale);
    const uint32_t zero_addr = smem_u32(smem + OFF_ZERO);
    const uint32_t so_ld = smem_u32(so + dim0);
    float *so_st = so + dim0 + ((g & 8) ? 4 : 0) + ((g & 4) ? 2 : 0) + ((g & 2) ? 1 : 0);
    float *pl_st = spl + warp * PLS + g;
    const float *pl_ld = spl + tid;
    const int jb = 2 * warp + hh;            // slice: thread tid < 128 is slot tid = g + 16 jb
    const float clip = c.clip;
    // With at most 32 slots a warp holds one slot per lane (slot = lane = g + 16 hh): every warp finishes all
    // logits itself and goes straight to the argmax - no second barrier, no exchange of the processed logits.
    // The partial logits then alternate between two buffers (a fast warp may already write step t + 1).
    constexpr bool FUSED = MM <= RL4CO_TSPG_FUSED_ROWS;
    constexpr int NS = FUSED ? (MM + 1) / 2 : 1;      // slots per lane in the fused finish: lane + 32 k = g + 16 (hh + 2 k)
    unsigned fkey[NS];
#pragma unroll
    for (int k = 0; k < NS; ++k)
        fkey[k] = ((unsigned)(lane + 32 * k) << 8) | (unsigned)(reinterpret_cast<const int *>(smem + OFF_NODE)[warp * 128 + lane + 32 * k] & 0xff);

    while (c.t < t_stop) {
        TG_TRACE_AT(0);
        // sampling = arg-max of logits + Gumbel noise (rollout_util.cuh), read from the ring
        float G[4] = {0.f, 0.f, 0.f, 0.f};
        if (SAMPLING) {
            if ((c.t & 7) == 0 && c.t > 0) gen_noise(c, c.t + 8 + warp, smem);
            const float4 g4 = reinterpret_cast<const float4 *>(smem + OFF_NOISE)[(c.t & 15) * 32 + lane];
            G[0] = g4.x; G[1] = g4.y; G[2] = g4.z; G[3] = g4.w;
        }
        // ---- query of this half head (decoder.py:141-142), pre-scaled; step 0 has no context row
        u64 q2[4];
        {
            u64 cr[4];
            lds_2x128(c.t == 0 ? zero_addr : c.c_addr + (uint32_t)c.cur * (E * 4), cr);
#pragma unroll
            for (int i = 0; i < 4; ++i) q2[i] = ffma2(cr[i], scale2, c.qbs2[i]);
        }
        // ---- glimpse scores of the own slots (attention.py:289-306), base-2 scaled
        float s[MM];
        {
            u64 acc[MM];
#pragma unroll
            for (int j = 0; j < MM; ++j) acc[j] = 0ull;
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < MM; ++j) acc[j] = ffma2(K2[j][i], q2[i], acc[j]);
#pragma unroll
            for (int j = 0; j < MM; ++j) s[j] = hsum2(acc[j]);
#pragma unroll
            for (int j = 0; j < MM; ++j) s[j] += __shfl_xor_sync(kFull, s[j], 16);
        }
        float mx = -INFINITY;
#pragma unroll
        for (int j = 0; j < MM; ++j) {
            s[j] = ((c.alive >> j) & 1u) ? s[j] : -INFINITY;
            mx = fmaxf(mx, s[j]);
        }
        mx = redux_max(mx);
        TG_TRACE_AT(1);
        float psum = 0.f;
        u64 oacc[4] = {0ull, 0ull, 0ull, 0ull};
#pragma unroll
        for (int j = 0; j < MM; ++j) {
            const float p = ex2f(s[j] - mx);        // visited: 2^-inf = 0
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
            *so_st = w1;                            // un-normalised; 1/psum is applied to the partial logits
        }
        __syncwarp();
        TG_TRACE_AT(2);
        // ---- partial pointer logits of this head, project_out folded (attention.py:276-282)
        {
            u64 o2[4];
            lds_2x128(so_ld, o2);
            u64 acc[MM];
#pragma unroll
            for (int j = 0; j < MM; ++j) acc[j] = 0ull;
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < MM; ++j) acc[j] = ffma2(L2[j][i], o2[i], acc[j]);
            const float inv = __fdividef(1.f, psum);
            float pl[MM];
#pragma unroll
            for (int j = 0; j < MM; ++j) pl[j] = hsum2(acc[j]) * inv;
#pragma unroll
            for (int j = 0; j < MM; ++j) pl[j] += __shfl_xor_sync(kFull, pl[j], 16);
            if (hh == 0) {
                float *dst = pl_st + ((FUSED && (c.t & 1)) ? kWarps * PLS : 0);
#pragma unroll
                for (int j = 0; j < MM; ++j) dst[16 * j] = pl[j];
            }
        }
        TG_TRACE_AT(3);
        __syncthreads();
        TG_TRACE_AT(4);
        float gm;
        unsigned win;
        if (FUSED) {
            const float *pl = spl + ((c.t & 1) ? kWarps * PLS : 0) + lane;
            float xv[NS];
#pragma unroll
            for (int k = 0; k < NS; ++k) {
                const float *pk = pl + 32 * k;
                const float sum = ((pk[0] + pk[2 * PLS]) + (pk[4 * PLS] + pk[6 * PLS])) +
                                  ((pk[PLS] + pk[3 * PLS]) + (pk[5 * PLS] + pk[7 * PLS]));
                const float e2 = ex2f(sum * (kInvSqrtE * 2.f * kLog2e));
                const float th = fmaf(-2.f * clip, rcpf(e2 + 1.f), clip);
                if (((c.own >> (hh + 2 * k)) & 1u) && sum != sum) c.status |= RL4CO_ST_NAN_LOGITS;
                xv[k] = ((c.alive >> (hh + 2 * k)) & 1u) ? th : -INFINITY;
            }
            TG_TRACE_AT(5);
            TG_TRACE_AT(6);
            float lm = xv[0];
#pragma unroll
            for (int k = 1; k < NS; ++k) lm = fmaxf(lm, xv[k]);
            gm = SAMPLING ? 0.f : redux_max(lm);    // max logit (sampling: the epilogue finds the row maxima itself)
            float yv[NS], ym = gm;
#pragma unroll
            for (int k = 0; k < NS; ++k) yv[k] = xv[k];
            if (SAMPLING) {
                ym = -INFINITY;
#pragma unroll
                for (int k = 0; k < NS; ++k) { yv[k] = xv[k] + G[k]; ym = fmaxf(ym, yv[k]); }
                ym = redux_max(ym);
            }
            unsigned cand = 0x7fffffffu;
#pragma unroll
            for (int k = NS - 1; k >= 0; --k) cand = yv[k] == ym ? fkey[k] : cand;
            win = redux_min_u32(cand);
            // one warp keeps the step's logits for the epilogue (the other slots of the row were set to -inf on entry)
            if (warp == (c.t & (kWarps - 1))) {
#pragma unroll
                for (int k = 0; k < NS; ++k)
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(c.x_addr + 4u * (lane + 32 * k)), "f"(xv[k]) : "memory");
            }
        } else {
        // ---- threads 0-127 = slots: logit = sum_h pl_h / sqrt(E); tanh clipping, mask (utils/decoding.py:141-191)
        if (tid < 128) {
            float xv = -INFINITY;
            if (jb < MM) {
                const float sum = ((pl_ld[0] + pl_ld[2 * PLS]) + (pl_ld[4 * PLS] + pl_ld[6 * PLS])) +
                                  ((pl_ld[PLS] + pl_ld[3 * PLS]) + (pl_ld[5 * PLS] + pl_ld[7 * PLS]));
                // clip * tanh(z) = clip - 2 clip / (2^(2 log2(e) z) + 1), z = sum / sqrt(E)
                const float e2 = ex2f(sum * (kInvSqrtE * 2.f * kLog2e));
                const float th = fmaf(-2.f * clip, rcpf(e2 + 1.f), clip);
                if (((c.own >> jb) & 1u) && sum != sum) c.status |= RL4CO_ST_NAN_LOGITS;
                xv = ((c.alive >> jb) & 1u) ? th : -INFINITY;
            }
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(c.x_addr + 4u * tid), "f"(xv) : "memory");
        }
        TG_TRACE_AT(5);
        __syncthreads();
        TG_TRACE_AT(6);
        // ---- every warp: all logits, lane = slots 4 lane .. 4 lane + 3; argmax, lowest slot on ties
        float x0, x1, x2, x3;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x0), "=f"(x1), "=f"(x2), "=f"(x3) : "r"(c.x_addr + 16u * lane));
        gm = SAMPLING ? 0.f : redux_max(fmaxf(fmaxf(x0, x1), fmaxf(x2, x3)));
        float ym = gm;
        if (SAMPLING) {
            x0 += G[0]; x1 += G[1]; x2 += G[2]; x3 += G[3];
            ym = redux_max(fmaxf(fmaxf(x0, x1), fmaxf(x2, x3)));
        }
        unsigned cand = 0x7fffffffu;
        cand = x3 == ym ? c.candk[3] : cand;
        cand = x2 == ym ? c.candk[2] : cand;
        cand = x1 == ym ? c.candk[1] : cand;
        cand = x0 == ym ? c.candk[0] : cand;
        win = redux_min_u32(cand);
        }
        if (win == 0x7fffffffu) win = reinterpret_cast<const int *>(smem + OFF_NODE)[warp * 128];   // NaN logits: slot 0
        const int slot = (int)(win >> 8), node = (int)(win & 0xffu);
        if (tid == 0) { s_act[c.t] = node; s_slot[c.t] = slot; if (!SAMPLING) s_mw[c.t] = gm; }
        // ---- transition (tsp/env.py:67-93); the slot's owner lanes check feasibility (utils/decoding.py:377-379)
        {
            const unsigned bit = (g == (slot & 15)) ? (1u << (slot >> 4)) : 0u;
            if (bit && !(c.alive & bit)) c.status |= RL4CO_ST_INFEASIBLE_ACTION;
            c.alive &= ~bit;
        }
        c.cur = node;                               // (step 0: the caller sets the first node's context half, below)
        TG_TRACE_AT(7);
        ++c.t;
        c.x_addr += 128 * 4;
    }
}

// ---------------------------------------------------------------------------------------------
// re-pack the surviving slots of this warp's head into the first MN rows, in slot (= node) order
// ---------------------------------------------------------------------------------------------
template <int MO, int MN>
__device__ __forceinline__ void compact(Ctx &c, const u64 (&Ko)[MO][4], const u64 (&Vo)[MO][4], const u64 (&Lo)[MO][4],
                                        u64 (&Kn)[MN][4], u64 (&Vn)[MN][4], u64 (&Ln)[MN][4], unsigned char *smem) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int hh = lane >> 4, g = lane & 15;
    int *node_w = reinterpret_cast<int *>(smem + OFF_NODE) + warp * 128;
    const uint32_t scr = smem_u32(smem + OFF_SCR) + (uint32_t)warp * (kScrSlots * 64);
    const int trow = c.trow + (MN == 4 ? 0 : 1);
    TG_TRACE_ROW(trow, 0);
    int rank[MO], nid[MO];
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < MO; ++j) {
        const unsigned bal = __ballot_sync(kFull, (c.alive >> j) & 1u) & 0xffffu;
        rank[j] = cnt + __popc(bal & ((1u << g) - 1u));
        cnt += __popc(bal);
        nid[j] = node_w[g + 16 * j];
    }
    __syncwarp();
    if (hh == 0) {
#pragma unroll
        for (int j = 0; j < MO; ++j)
            if ((c.alive >> j) & 1u) node_w[rank[j]] = nid[j];
    }
    // scratch layout [half head][16-byte piece][slot]: consecutive slots are consecutive 16-byte words,
    // so both the rank-addressed stores and the slot-addressed loads are free of bank conflicts
    const uint32_t scr_h = scr + (uint32_t)hh * (2u * kScrSlots * 16u);
    auto move = [&](const u64 (&To)[MO][4], u64 (&Tn)[MN][4]) {
#pragma unroll
        for (int j = 0; j < MO; ++j)
            if (((c.alive >> j) & 1u) && rank[j] < 16 * MN) {
                const uint32_t ad = scr_h + (uint32_t)rank[j] * 16u;
                asm volatile("st.shared.v2.u64 [%0], {%1, %2};" ::"r"(ad), "l"(To[j][0]), "l"(To[j][1]) : "memory");
                asm volatile("st.shared.v2.u64 [%0], {%1, %2};" ::"r"(ad + kScrSlots * 16u), "l"(To[j][2]), "l"(To[j][3]) : "memory");
            }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < MN; ++j) {
            const int slot = g + 16 * j;
            if (slot < cnt) {
                const uint32_t ad = scr_h + (uint32_t)slot * 16u;
                asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(Tn[j][0]), "=l"(Tn[j][1]) : "r"(ad));
                asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(Tn[j][2]), "=l"(Tn[j][3]) : "r"(ad + kScrSlots * 16u));
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) Tn[j][i] = 0ull;
            }
        }
        __syncwarp();
    };
    TG_TRACE_ROW(trow, 1);
    move(Ko, Kn);
    TG_TRACE_ROW(trow, 2);
    move(Vo, Vn);
    TG_TRACE_ROW(trow, 3);
    move(Lo, Ln);
    TG_TRACE_ROW(trow, 4);
    unsigned m = 0;
#pragma unroll
    for (int j = 0; j < MN; ++j) m |= (g + 16 * j < cnt) ? (1u << j) : 0u;
    c.alive = c.own = m;
#pragma unroll
    for (int k = 0; k < 4; ++k) c.candk[k] = ((unsigned)(4 * lane + k) << 8) | (unsigned)(node_w[4 * lane + k] & 0xff);
}

// MS: POMO multistart (work item = rollout, forced first action).  A compile-time switch because the extra live values
// of the multistart bookkeeping cost the single-start step loop 4 % (8.22 vs 7.90 ms per 16 280 TSP100 rollouts) when they
// are run-time state of the same kernel.
template <int M0, bool SAMPLING, bool MS>
__global__ void __launch_bounds__(kThreads, 1)
am_rollout_tsp_greedy_kernel(const rl4co_am_rollout_args a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int N = a.N, T = a.T;
    float *cbuf = reinterpret_cast<float *>(smem + off_c(T));
    float *sxs = reinterpret_cast<float *>(smem + OFF_XS);
    float *s_mw = reinterpret_cast<float *>(smem + OFF_MW);
    int *s_act = reinterpret_cast<int *>(smem + OFF_ACT);
    int *s_slot = reinterpret_cast<int *>(smem + OFF_SLOT);
    float *s_lp = reinterpret_cast<float *>(smem + OFF_LP);
    float2 *s_locs = reinterpret_cast<float2 *>(smem + OFF_LOCS);
    float *spart = reinterpret_cast<float *>(smem + OFF_PART);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + OFF_BAR);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int hh = lane >> 4, g = lane & 15;
    const int dim0 = 16 * warp + 8 * hh;
    const int64_t Mrows = a.B * (int64_t)N;
    const size_t c_bytes = (size_t)N * E * 4;

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_barrier_init();
    }
    if (tid < E) reinterpret_cast<float *>(smem + OFF_ZERO)[tid] = 0.f;
    __syncthreads();
    unsigned par0 = 0u, par1 = 0u;
    auto stage_context = [&](int64_t b, int buf) {  // thread 0: TMA the context table of instance b
        mbar_expect_tx(&bar[buf], (unsigned)c_bytes);
        const char *src = reinterpret_cast<const char *>(a.tables + ((size_t)3 * Mrows + b * (int64_t)N) * E);
        char *dst = reinterpret_cast<char *>(cbuf + (size_t)buf * N * E);
        for (size_t off = 0; off < c_bytes; off += 16384)
            tma_load_1d(dst + off, src + off, (unsigned)(c_bytes - off < 16384 ? c_bytes - off : 16384), &bar[buf]);
    };
    // work item = one rollout (instance b, start s), flat index r = s * B + b (utils/ops.py:11-34).  Consecutive items go to
    // different CTAs, so the ~1.5 instances whose starts are in flight at any time keep their tables hot in L2.
    const int S = MS ? (a.S > 1 ? a.S : 1) : 1;
    const int64_t W = MS ? a.B * (int64_t)S : a.B;
    constexpr int t_first = MS ? 1 : 0;             // multistart: step 0 is the forced POMO start node
    // table 4 left out (a.w_first): half of its weight block waits in the warp's compaction scratch (4 KB, dims 4..7 of every
    // thread), which nothing else uses before the first compaction.  (Sending the instance's block of h to L2 one instance
    // ahead, so that the first node's row is an L2 hit at the end of step 0, was measured to change nothing - step 0 takes
    // 3.7 k clk either way - and cost 51 KB of DRAM reads per rollout; not done.)
    float *fc_scr = reinterpret_cast<float *>(smem + OFF_SCR) + warp * (kScrSlots * 16);
    if (tid == 0 && (int64_t)blockIdx.x < W) stage_context(MS ? blockIdx.x / S : blockIdx.x, 0);

    int it = 0;
    for (int64_t w = blockIdx.x; w < W; w += gridDim.x, ++it) {
        const int64_t b = MS ? w / S : w;
        const int s = MS ? (int)(w % S) : 0;
        const int64_t r = MS ? (int64_t)s * a.B + b : w;
        const int buf = it & 1;
#ifdef RF_TRACE
        Ctx c;
        c.trace = (blockIdx.x == 0 && it == 1) ? (void *)a.logp_all : nullptr;
        TG_TRACE_ROW(T + 2, 0);
#endif
        // ------------------------------------------------ registers <- tables (global / L2)
        u64 K2[M0][4], V2[M0][4], L2[M0][4];
        {
            const float *kb = a.tables + ((size_t)0 * Mrows + b * (int64_t)N) * E + dim0;
            const float *vb = a.tables + ((size_t)1 * Mrows + b * (int64_t)N) * E + dim0;
            const float *lb = a.tables + ((size_t)2 * Mrows + b * (int64_t)N) * E + dim0;
#pragma unroll
            for (int j = 0; j < M0; ++j) {
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
        }
#ifndef RF_TRACE
        Ctx c;
#endif
        // constant part of the query, pre-scaled: scale * (gctx + extra); extra = step-0 placeholder
        // (context.py:111-121), then the first node's half of the context projection
        auto set_qb_vals = [&](const float (&x)[8]) {
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = 0.f;
            if (a.gctx) {
                const float4 g0 = __ldg(reinterpret_cast<const float4 *>(a.gctx + b * E + dim0));
                const float4 g1 = __ldg(reinterpret_cast<const float4 *>(a.gctx + b * E + dim0) + 1);
                v[0] = g0.x; v[1] = g0.y; v[2] = g0.z; v[3] = g0.w; v[4] = g1.x; v[5] = g1.y; v[6] = g1.z; v[7] = g1.w;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] += x[i];
#pragma unroll
            for (int i = 0; i < 4; ++i) c.qbs2[i] = pack2(v[2 * i] * kScoreScale, v[2 * i + 1] * kScoreScale);
        };
        auto set_qb = [&](const float *extra) {
            const float4 e0 = __ldg(reinterpret_cast<const float4 *>(extra));
            const float4 e1 = __ldg(reinterpret_cast<const float4 *>(extra) + 1);
            const float x[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
            set_qb_vals(x);
        };
        set_qb(a.q0 + dim0);
        // row `node` of table 4, or - when the caller left that table out - formed here from h and w_first
        auto set_first = [&](int node) {
            if (a.w_first) {        // the warp's slice of the glimpse buffer is free between two steps, its compaction
                                    // scratch before the first compaction
                float *o16 = reinterpret_cast<float *>(smem + OFF_O) + 16 * warp;
                first_context_row(a.h + (b * (int64_t)N + node) * E, a.w_first, o16, fc_scr, true, dim0, g);
                const float4 e0 = *reinterpret_cast<const float4 *>(o16 + 8 * hh);
                const float4 e1 = *reinterpret_cast<const float4 *>(o16 + 8 * hh + 4);
                const float x[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
                set_qb_vals(x);
                __syncwarp();
            } else {
                set_qb(a.tables + ((size_t)4 * Mrows + b * (int64_t)N + node) * E + dim0);
            }
        };

        TG_TRACE_ROW(T + 2, 1);
        // ------------------------------------------------ small per-instance state -> smem
        __syncthreads();                            // previous instance's epilogue is done with smem
        {
            const float2 *lrow = reinterpret_cast<const float2 *>(a.locs + b * (int64_t)N * 2);
            for (int n = tid; n < N; n += kThreads) s_locs[n] = lrow[n];
            int *node_w = reinterpret_cast<int *>(smem + OFF_NODE) + warp * 128;
#pragma unroll
            for (int k = 0; k < 4; ++k) node_w[4 * lane + k] = 4 * lane + k;
        }
        // table 4 left out: half of its weight block starts towards the warp's compaction scratch now (free until the first
        // compaction) and lands during step 0
        if (a.w_first) first_context_prefetch(a.w_first, fc_scr, dim0, g);
        if (SAMPLING) {                             // Gumbel noise of steps 0..15 (warp w: steps w and w + 8)
            c.seed = a.seed; c.offset = a.offset; c.r = (uint64_t)r;
            gen_noise(c, warp, smem);
            gen_noise(c, warp + 8, smem);
        }
        if (tid == 0 && w + gridDim.x < W) {        // next item: context table -> other buffer, K/V/L rows -> L2
            const int64_t nb = MS ? (w + gridDim.x) / S : w + gridDim.x;
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            stage_context(nb, buf ^ 1);
            for (int tb = 0; tb < 3; ++tb) {
                const char *src = reinterpret_cast<const char *>(a.tables + ((size_t)tb * Mrows + nb * (int64_t)N) * E);
                for (size_t off = 0; off < c_bytes; off += 32768)
                    prefetch_l2(src + off, (unsigned)(c_bytes - off < 32768 ? c_bytes - off : 32768));
            }
        }
        __syncthreads();

        TG_TRACE_ROW(T + 2, 2);
        // ------------------------------------------------ env state (tsp/env.py:95-120)
        c.own = 0;
#pragma unroll
        for (int j = 0; j < M0; ++j) c.own |= (g + 16 * j < N) ? (1u << j) : 0u;
        c.alive = c.own;
#pragma unroll
        for (int k = 0; k < 4; ++k) c.candk[k] = (unsigned)(4 * lane + k) * 257u;      // slot == node
        c.cur = 0;
        c.t = 0;
        c.status = 0;
        c.clip = a.tanh_clipping;
        c.seed = a.seed; c.offset = a.offset; c.r = (uint64_t)r;
        c.c_addr = smem_u32(cbuf + (size_t)buf * N * E + dim0);
        c.x_addr = smem_u32(sxs);
#ifdef RF_TRACE
        c.trace = (blockIdx.x == 0 && it == 1) ? (void *)a.logp_all : nullptr;
#else
        c.trace = nullptr;
#endif
        c.trow = T;
        const int t_end = N < T ? N : T;

        if (buf == 0) { mbar_wait(&bar[0], par0); par0 ^= 1u; }      // context table landed?
        else          { mbar_wait(&bar[1], par1); par1 ^= 1u; }

        if (MS) {                                   // forced POMO start node, log-prob 0 (utils/decoding.py:287-305)
            const int a0 = s % a.num_loc;           // slot == node before the first compaction
            c.alive &= ~((g == (a0 & 15)) ? (1u << (a0 >> 4)) : 0u);
            if (tid == 0) { s_act[0] = a0; s_slot[0] = a0; s_lp[0] = 0.f; }
            set_first(a0);
            c.cur = a0;
            c.t = 1;
            c.x_addr += 128 * 4;
        }
        TG_TRACE_ROW(T + 2, 3);
        // fused stages keep only the live slots of a step's logits: the rest of the rows is -inf from here on
        auto fill_rows = [&](int t0) {
            for (int i = tid; i < (t_end - t0) * 128; i += kThreads) sxs[t0 * 128 + i] = -INFINITY;
        };
        // ================================================== decode, in stages of fewer and fewer rows
        // after t steps N - t nodes are unvisited: 4 rows hold them from t = N - 64 on, 2 rows from t = N - 32
        // ctx = Wc [h_first ; h_cur] from step 1 on (context.py:123-131): the stage that holds step 0 is entered twice, for
        // step 0 and for the rest, and the first node's half is set in between - OUTSIDE the step loop.  With table 4 left
        // out that is a call and 18 loads; inside the loop body they changed its register allocation and cost every step
        // (K2 509 -> 533 ms per 2^20 rollouts even when the branch was never taken).
        auto with_first = [&](auto &&run, int stop) {
#pragma unroll 1
            for (int pass = (c.t == 0 && stop > 0) ? 0 : 1; pass < 2; ++pass) {
                run(pass == 0 ? 1 : stop);
                if (pass == 0) set_first(c.cur);
            }
        };
        if (M0 > 4) {
            const int stop = N - 64 < t_end ? N - 64 : t_end;
            with_first([&](int t_stop) { run_stage<M0, SAMPLING>(c, K2, V2, L2, t_stop, smem); }, stop);
        }
        if (M0 > 4 ? c.t < t_end : true) {
            u64 K4[4][4], V4[4][4], L4[4][4];
            if (M0 > 4) {
                compact<M0, 4>(c, K2, V2, L2, K4, V4, L4, smem);
            } else if (M0 == 4) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) { K4[j][i] = K2[j < M0 ? j : 0][i]; V4[j][i] = V2[j < M0 ? j : 0][i]; L4[j][i] = L2[j < M0 ? j : 0][i]; }
            }
            if (M0 >= 4) {
                const int stop = N - 32 < t_end ? N - 32 : t_end;
                if (RL4CO_TSPG_FUSED_ROWS >= 4) fill_rows(c.t);
                with_first([&](int t_stop) { run_stage<4, SAMPLING>(c, K4, V4, L4, t_stop, smem); }, stop);
            }
            if (c.t < t_end) {
                u64 Ka[2][4], Va[2][4], La[2][4];
                if (M0 >= 4) {
                    compact<4, 2>(c, K4, V4, L4, Ka, Va, La, smem);
                } else {
#pragma unroll
                    for (int j = 0; j < 2; ++j)
#pragma unroll
                        for (int i = 0; i < 4; ++i) { Ka[j][i] = K2[j < M0 ? j : 0][i]; Va[j][i] = V2[j < M0 ? j : 0][i]; La[j][i] = L2[j < M0 ? j : 0][i]; }
                }
                if (RL4CO_TSPG_FUSED_ROWS < 4 || M0 < 4) fill_rows(c.t);
                with_first([&](int t_stop) { run_stage<2, SAMPLING>(c, Ka, Va, La, t_stop, smem); }, t_end);
            }
        }

        TG_TRACE_ROW(T + 2, 4);
        // ================================================== epilogue =====================
        __syncthreads();
        const int nact = c.t;
        int status = c.status;
        // chosen log-probs: logp = (x_a - m) - log(sum exp(x - m))   (utils/decoding.py:191,364-368); the
        // exponentials through ex2.approx (relative error 2^-22; masked logits: 2^-inf = 0)
        {
            // warp w takes steps w, w + 8, ...: all of its (at most 16) sums are reduced together, then lane u
            // finishes step w + 8 u, so neither the shuffles nor the logarithm are serialised over the steps
            constexpr int U = kMaxT / kWarps;
            float es[U], mm[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int q = warp + u * kWarps;
                es[u] = 0.f;
                mm[u] = 0.f;
                if (q >= t_first && q < nact) {
                    const float4 xv = *reinterpret_cast<const float4 *>(sxs + q * 128 + 4 * lane);
                    mm[u] = SAMPLING ? redux_max(fmaxf(fmaxf(xv.x, xv.y), fmaxf(xv.z, xv.w))) : s_mw[q];
                    es[u] = (ex2f((xv.x - mm[u]) * kLog2e) + ex2f((xv.y - mm[u]) * kLog2e)) +
                            (ex2f((xv.z - mm[u]) * kLog2e) + ex2f((xv.w - mm[u]) * kLog2e));
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
#pragma unroll
                for (int u = 0; u < U; ++u) es[u] += __shfl_xor_sync(kFull, es[u], o);
            float my_es = 1.f, my_m = 0.f;
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (lane == u) { my_es = es[u]; my_m = mm[u]; }
            const int q = warp + lane * kWarps;
            if (lane < U && q >= t_first && q < nact) {
                const float lpa = (sxs[q * 128 + s_slot[q]] - my_m) - logf(my_es);
                s_lp[q] = lpa;
                if (!(lpa > -1000.f)) status |= RL4CO_ST_NEG_INF_LOGP;
            }
        }
        TG_TRACE_ROW(T + 2, 5);
        float part = 0.f;
        for (int q = tid; q < nact; q += kThreads) {     // tsp/env.py:157-163
            float2 p = s_locs[s_act[q]], w = s_locs[s_act[q + 1 == nact ? 0 : q + 1]];
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
                if (a.steps) a.steps[r] = nact;
                if (st) raise_status(a.status, st);
            }
        }
        for (int q = tid; q < T; q += kThreads) {
            a.actions[(size_t)r * T + q] = q < nact ? (int64_t)s_act[q] : 0;
            if (a.logp) a.logp[(size_t)r * T + q] = q < nact ? s_lp[q] : 0.f;
        }
        TG_TRACE_ROW(T + 2, 6);
    }
}

template <int M0, bool SAMPLING, bool MS>
static int launch(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    const size_t total = smem_total(a.N, a.T);
    cudaError_t e = cudaFuncSetAttribute(am_rollout_tsp_greedy_kernel<M0, SAMPLING, MS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)total);
    if (e != cudaSuccess) {
        set_error("am_rollout(tsp greedy): cannot reserve %zu B of shared memory: %s", total, cudaGetErrorString(e));
        return 2;
    }
    const int64_t W = MS ? a.B * (int64_t)(a.S > 1 ? a.S : 1) : a.B;
    const int64_t grid = W < sm_count ? W : sm_count;       // persistent: one CTA per SM
    am_rollout_tsp_greedy_kernel<M0, SAMPLING, MS><<<(unsigned)grid, kThreads, total, stream>>>(a);
    RL4CO_CHECK_LAUNCH("am_rollout(tsp greedy)");
    return 0;
}

}  // namespace tspg

bool rollout_tsp_greedy_supports(const rl4co_am_rollout_args &a) {
#ifndef RF_TRACE
    if (a.logp_all != nullptr) return false;
#endif
    if (a.env != RL4CO_ENV_TSP || a.mode == RL4CO_MODE_EVALUATE) return false;
    if (a.S > 1 && !a.multistart) return false;      // several free starts per instance: the generic kernel
    if (a.start_nodes != nullptr) return false;      // caller-chosen start nodes: the generic kernel
    if (!(a.tanh_clipping > 0.f) || a.temperature != 1.f || a.mask_logits == 0) return false;
    if (a.N > 112 || a.T > tspg::kMaxT || a.N < 2) return false;
    return tspg::smem_total(a.N, a.T) <= (size_t)227 * 1024;
}

template <bool SAMPLING, bool MS>
static int launch_rows(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    if (a.N <= 32) return tspg::launch<2, SAMPLING, MS>(a, stream);
    if (a.N <= 64) return tspg::launch<4, SAMPLING, MS>(a, stream);
    return tspg::launch<7, SAMPLING, MS>(a, stream);
}

int launch_rollout_tsp_greedy(const rl4co_am_rollout_args &a, cudaStream_t stream) {
    const bool ms = a.multistart != 0;
    if (a.mode == RL4CO_MODE_SAMPLING) return ms ? launch_rows<true, true>(a, stream) : launch_rows<true, false>(a, stream);
    return ms ? launch_rows<false, true>(a, stream) : launch_rows<false, false>(a, stream);
}

}  // namespace rl4co
