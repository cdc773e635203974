// K2 - persistent Attention-Model rollout kernel (decode + environment in one launch).
//
// One CTA per rollout r = s*B + b.  The instance's per-node tables (glimpse K, glimpse V,
// folded logit K, per-node context) are staged ONCE into shared memory with TMA bulk copies
// (cp.async.bulk + mbarrier complete_tx), rows padded to 132 floats so that the
// "8 consecutive nodes x one 16-byte chunk" access of a quarter warp is bank-conflict free.
// The whole trajectory then runs inside the kernel: context query, 8-head masked glimpse,
// folded pointer logits, tanh clipping, masked log-softmax, greedy / Philox sampling /
// evaluate selection, the TSP / CVRP / SLAP transition + action mask, and at the end the
// reward.  State (visited / feasible bytes, capacity, assignment) never leaves the SM.
//
// Reference op order being reproduced (all fp32): SURVEY.md appendix A;
// rl4co/models/zoo/am/decoder.py:134-200, rl4co/models/nn/attention.py:263-309,
// rl4co/utils/decoding.py:141-191,327-397, rl4co/envs/routing/{tsp,cvrp}/env.py,
// rl4co/envs/warehousing/slap/env.py:38-143.
#include "rollout_util.cuh"

namespace rl4co {

constexpr int E = RL4CO_EMBED_DIM;      // 128
constexpr int H = RL4CO_NUM_HEADS;      // 8
constexpr int LD = E + 4;               // padded row stride (floats)
constexpr int kThreads = 512;           // 16 warps: thread = (node slot 0..127, 32-dim chunk 0..3)
constexpr int kWarps = kThreads / 32;
constexpr int kMaxNodes = 128;          // node slots of the thread mapping
constexpr int SA_LD = kMaxNodes + 1;    // attention weights stored [head][node], odd stride

// ---------------------------------------------------------------- smem layout ------
struct Layout {
    int tab_rows;       // N
    int n_smem_tab;     // tables resident in shared memory (3 or 4)
    size_t off_tab, off_q, off_o, off_base, off_a, off_part, off_x, off_feas, off_vis, off_act,
        off_lp, off_locs, off_dem, off_misc, off_bar, total;
};

__host__ __device__ inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

__host__ __device__ inline Layout make_layout(int N, int T, int n_smem_tab) {
    Layout L;
    L.tab_rows = N;
    L.n_smem_tab = n_smem_tab;
    size_t o = 0;
    L.off_tab = o;  o += (size_t)n_smem_tab * N * LD * 4;
    L.off_q = o;    o += E * 4;
    L.off_o = o;    o += E * 4;
    L.off_base = o; o += E * 4;
    L.off_a = o;    o += align_up((size_t)H * SA_LD * 4, 16);
    L.off_part = o; o += (size_t)kWarps * E * 4;
    L.off_x = o;    o += kMaxNodes * 4;
    L.off_feas = o; o += kMaxNodes;
    L.off_vis = o;  o += kMaxNodes;
    L.off_act = o;  o += align_up((size_t)T * 4, 16);
    L.off_lp = o;   o += align_up((size_t)T * 4, 16);
    L.off_locs = o; o += align_up((size_t)N * 8, 16);
    L.off_dem = o;  o += align_up((size_t)N * 4, 16);
    L.off_misc = o; o += 64;
    L.off_bar = o;  o += 16;
    L.total = o;
    return L;
}

// misc words
enum { M_ACTION = 0, M_CUR = 1, M_FIRST = 2, M_DONE = 3, M_STATUS = 4, M_STEP = 5, M_USED = 6, M_NACT = 7 };

// dot product of 16 consecutive floats (4 x float4), sequential fma chain
__device__ __forceinline__ float dot16(const float4 *a, const float4 *b) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float4 x = a[k], y = b[k];
        s = fmaf(x.x, y.x, s); s = fmaf(x.y, y.y, s); s = fmaf(x.z, y.z, s); s = fmaf(x.w, y.w, s);
    }
    return s;
}

template <int ENV>
__global__ void __launch_bounds__(kThreads, 1)
am_rollout_kernel(const rl4co_am_rollout_args a, const int n_smem_tab, const int64_t R,
                  const int64_t chunk) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int N = a.N, T = a.T;
    const Layout L = make_layout(N, T, n_smem_tab);
    float *Ks = reinterpret_cast<float *>(smem + L.off_tab);
    float *Vs = Ks + (size_t)N * LD;
    float *Ls = Vs + (size_t)N * LD;
    float *Cs = n_smem_tab > 3 ? Ls + (size_t)N * LD : nullptr;
    float *sq = reinterpret_cast<float *>(smem + L.off_q);
    float *so = reinterpret_cast<float *>(smem + L.off_o);
    float *sbase = reinterpret_cast<float *>(smem + L.off_base);
    float *sa = reinterpret_cast<float *>(smem + L.off_a);
    float *spart = reinterpret_cast<float *>(smem + L.off_part);
    float *sx = reinterpret_cast<float *>(smem + L.off_x);
    uint8_t *feas = smem + L.off_feas;
    uint8_t *vis = smem + L.off_vis;
    int *s_act = reinterpret_cast<int *>(smem + L.off_act);
    float *s_lp = reinterpret_cast<float *>(smem + L.off_lp);
    float2 *s_locs = reinterpret_cast<float2 *>(smem + L.off_locs);
    float *s_dem = reinterpret_cast<float *>(smem + L.off_dem);
    int *misc = reinterpret_cast<int *>(smem + L.off_misc);
    float *miscf = reinterpret_cast<float *>(smem + L.off_misc);
    uint64_t *bar = reinterpret_cast<uint64_t *>(smem + L.off_bar);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t M = a.B * (int64_t)N;              // rows per table
    const int S = a.S > 1 ? a.S : 1;
    const float inv_temp_div = a.temperature;        // logits / temperature (true division)
    const float sqrt_e = sqrtf((float)E);

    if (tid == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    unsigned parity = 0;
    int64_t loaded_b = -1;

    const int64_t w_begin = blockIdx.x * chunk;
    const int64_t w_end = (w_begin + chunk < R) ? w_begin + chunk : R;
    for (int64_t w = w_begin; w < w_end; ++w) {
        // work item -> (instance b, start s); flat rollout index r = s*B + b (utils/ops.py:11-34)
        const int64_t b = w / S;
        const int s = (int)(w % S);
        const int64_t r = (int64_t)s * a.B + b;

        // ------------------------------------------------ stage tables (TMA) + small state
        if (b != loaded_b) {
            __syncthreads();                         // everyone is done with the previous tables
            if (warp == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                if (lane == 0) mbar_expect_tx(bar, (unsigned)(n_smem_tab * N * E * 4));
                __syncwarp();
                for (int t = 0; t < n_smem_tab; ++t) {
                    const float *src = a.tables + ((size_t)t * M + b * (int64_t)N) * E;
                    float *dst = Ks + (size_t)t * N * LD;
                    for (int n = lane; n < N; n += 32)
                        tma_load_1d(dst + (size_t)n * LD, src + (size_t)n * E, E * 4, bar);
                }
            }
            if (tid < E) sbase[tid] = a.gctx ? a.gctx[b * E + tid] : 0.f;
            const float2 *lrow = reinterpret_cast<const float2 *>(a.locs + b * (int64_t)N * 2);
            for (int n = tid; n < N; n += kThreads) s_locs[n] = lrow[n];
            if (ENV == RL4CO_ENV_CVRP)
                for (int n = tid; n < N - 1; n += kThreads) s_dem[n] = a.demand[b * (int64_t)(N - 1) + n];
        }
        // reset env state (tsp/env.py:95-120, cvrp/env.py:107-135, slap/env.py:95-129)
        for (int n = tid; n < kMaxNodes; n += kThreads) {
            vis[n] = 0;
            feas[n] = (n < N) && !((ENV != RL4CO_ENV_TSP) && n == 0);   // cvrp: at depot with customers left; slap: depot unavailable
        }
        if (tid == 0) {
            misc[M_CUR] = 0; misc[M_FIRST] = 0; misc[M_DONE] = 0; misc[M_STATUS] = 0;
            misc[M_STEP] = 0; misc[M_NACT] = 0; miscf[M_USED] = 0.f;
        }
        if (b != loaded_b) {
            mbar_wait(bar, parity);
            parity ^= 1;
            loaded_b = b;
        }
        __syncthreads();

        const float cap = (ENV == RL4CO_ENV_CVRP) ? a.vehicle_capacity[b] : 0.f;
        float ll = 0.f;          // running sum of chosen log-probs (warp 0, lane 0)
        float first_ctx = 0.f;   // tsp: first-node context row element (threads < E)
        bool have_first = false;

        // ---- CVRP action mask from (visited, used, current) - cvrp/env.py:137-149. warp 0.
        auto cvrp_mask = [&](int cur, float used) {
            int any = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int n = lane * 4 + k;
                if (n >= 1 && n < N) {
                    bool exceeds = __fadd_rn(s_dem[n - 1], used) > cap;
                    bool blocked = (vis[n] != 0) | exceeds;
                    feas[n] = blocked ? 0 : 1;
                    any |= !blocked;
                }
            }
            any = __any_sync(kFull, any);
            if (lane == 0) feas[0] = (cur == 0 && any) ? 0 : 1;
        };
        // ---- environment transition for the chosen action. warp 0, all lanes.
        auto transition = [&](int act) {
            const int istep = misc[M_STEP];
            __syncwarp();
            if (ENV == RL4CO_ENV_TSP) {                     // tsp/env.py:67-93
                if (lane == 0) {
                    feas[act] = 0; vis[act] = 1;
                    if (istep == 0) misc[M_FIRST] = act;
                    misc[M_CUR] = act; misc[M_STEP] = istep + 1;
                    misc[M_DONE] = (istep + 1 >= N);
                }
            } else if (ENV == RL4CO_ENV_CVRP) {             // cvrp/env.py:73-105
                float used = miscf[M_USED];
                int idx = act - 1 < 0 ? 0 : act - 1;
                used = __fmul_rn(__fadd_rn(used, s_dem[idx]), act != 0 ? 1.f : 0.f);
                __syncwarp();
                if (lane == 0) vis[act] = 1;
                __syncwarp();
                int cnt = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) { int n = lane * 4 + k; if (n < N) cnt += vis[n]; }
                cnt = warp_sum_int(cnt);
                cvrp_mask(act, used);
                if (lane == 0) {
                    miscf[M_USED] = used; misc[M_CUR] = act; misc[M_STEP] = istep + 1;
                    misc[M_DONE] = (cnt == N);
                }
            } else {                                        // slap/env.py:38-93
                if (lane == 0) {
                    feas[act] = 0; vis[act] = 1;
                    misc[M_CUR] = act; misc[M_STEP] = istep + 1;
                    misc[M_DONE] = (istep + 1 >= a.P);
                }
            }
            __syncwarp();
        };

        if (warp == 0) {
            if (ENV == RL4CO_ENV_CVRP) cvrp_mask(0, 0.f);
            if (a.multistart) {     // forced POMO start node, log-prob 0 (utils/decoding.py:287-305)
                int a0 = a.start_nodes ? (int)a.start_nodes[r] : s % a.num_loc + (ENV == RL4CO_ENV_TSP ? 0 : 1);
                if (lane == 0) { s_act[0] = a0; s_lp[0] = 0.f; misc[M_NACT] = 1; }
                if (a.logp_all)
                    for (int n = lane; n < N; n += 32) a.logp_all[((size_t)r * T) * N + n] = 0.f;
                transition(a0);
            }
        }
        __syncthreads();

        // ================================================== decode loop ==================
        while (!misc[M_DONE] && misc[M_NACT] < T) {
            const int cur = misc[M_CUR];
            const int istep = misc[M_STEP];
            const int t = misc[M_NACT];
            // ---- context query q = ctx + graph context (am/decoder.py:134-146)
            if (tid < E) {
                float ctx = 0.f;
                if (ENV == RL4CO_ENV_TSP) {                 // context.py:102-137 (upstream TSP context)
                    if (istep == 0) {
                        ctx = a.q0[tid];
                    } else {
                        if (!have_first) {
                            if (a.w_first) {    // table 4 left out (args h / w_first): its one row, formed here
                                const float4 *hr = reinterpret_cast<const float4 *>(a.h + (b * (int64_t)N + misc[M_FIRST]) * E);
                                const float4 *wr = reinterpret_cast<const float4 *>(a.w_first + (size_t)tid * E);
                                first_ctx = 0.f;
                                for (int k = 0; k < E / 4; ++k) {
                                    const float4 x = __ldg(hr + k), w = __ldg(wr + k);
                                    first_ctx = fmaf(x.x, w.x, first_ctx); first_ctx = fmaf(x.y, w.y, first_ctx);
                                    first_ctx = fmaf(x.z, w.z, first_ctx); first_ctx = fmaf(x.w, w.w, first_ctx);
                                }
                            } else {
                                first_ctx = a.tables[((size_t)4 * M + b * (int64_t)N + misc[M_FIRST]) * E + tid];
                            }
                            have_first = true;
                        }
                        float c = Cs ? Cs[(size_t)cur * LD + tid]
                                     : a.tables[((size_t)3 * M + b * (int64_t)N + cur) * E + tid];
                        ctx = first_ctx + c;
                    }
                } else if (ENV == RL4CO_ENV_CVRP) {         // context.py:58-71,317-331
                    float c = Cs ? Cs[(size_t)cur * LD + tid]
                                 : a.tables[((size_t)3 * M + b * (int64_t)N + cur) * E + tid];
                    float rem = __fsub_rn(cap, miscf[M_USED]);
                    ctx = fmaf(rem, a.wcap[tid], c);
                }
                sq[tid] = ctx + sbase[tid];
            }
            __syncthreads();

            // ---- glimpse scores: thread (node n, chunk c) -> heads 2c, 2c+1 (attention.py:289-306)
            const int c = lane >> 3, n = warp * 8 + (lane & 7);
            if (n < N) {
                float s0 = -INFINITY, s1 = -INFINITY;
                if (feas[n]) {
                    const float4 *kp = reinterpret_cast<const float4 *>(Ks + (size_t)n * LD + 32 * c);
                    const float4 *qp = reinterpret_cast<const float4 *>(sq + 32 * c);
                    s0 = dot16(kp, qp) * 0.25f;
                    s1 = dot16(kp + 4, qp + 4) * 0.25f;
                }
                sa[(2 * c) * SA_LD + n] = s0;
                sa[(2 * c + 1) * SA_LD + n] = s1;
            }
            __syncthreads();

            // ---- softmax over nodes, one warp per head
            if (warp < H) {
                float v[4], m = -INFINITY;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    int nn = lane + 32 * k;
                    v[k] = nn < N ? sa[warp * SA_LD + nn] : -INFINITY;
                    m = fmaxf(m, v[k]);
                }
                m = warp_max(m);
                float sum = 0.f;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    v[k] = (v[k] == -INFINITY) ? 0.f : expf(v[k] - m);
                    sum += v[k];
                }
                sum = warp_sum(sum);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    int nn = lane + 32 * k;
                    if (nn < N) sa[warp * SA_LD + nn] = v[k] / sum;
                }
            }
            __syncthreads();

            // ---- heads o = attn . V : warp g takes nodes g, g+16, ...; lane owns 4 dims
            {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                const int j = lane >> 2;
                for (int nn = warp; nn < N; nn += kWarps) {
                    if (!feas[nn]) continue;                 // masked nodes carry weight 0
                    float wgt = sa[j * SA_LD + nn];
                    float4 vv = *reinterpret_cast<const float4 *>(Vs + (size_t)nn * LD + 4 * lane);
                    acc.x = fmaf(wgt, vv.x, acc.x); acc.y = fmaf(wgt, vv.y, acc.y);
                    acc.z = fmaf(wgt, vv.z, acc.z); acc.w = fmaf(wgt, vv.w, acc.w);
                }
                *reinterpret_cast<float4 *>(spart + warp * E + 4 * lane) = acc;
            }
            __syncthreads();
            if (tid < E) {
                float sum = 0.f;
#pragma unroll
                for (int g = 0; g < kWarps; ++g) sum += spart[g * E + tid];
                so[tid] = sum;
            }
            __syncthreads();

            // ---- pointer logits with folded project_out (attention.py:276-285), then
            //      process_logits (utils/decoding.py:141-191): tanh clip, mask, temperature
            {
                float p = 0.f;
                const bool live = n < N && (feas[n] || !a.mask_logits);
                if (live) {
                    const float4 *lp4 = reinterpret_cast<const float4 *>(Ls + (size_t)n * LD + 32 * c);
                    const float4 *op4 = reinterpret_cast<const float4 *>(so + 32 * c);
                    p = dot16(lp4, op4) + dot16(lp4 + 4, op4 + 4);
                }
                p += __shfl_xor_sync(kFull, p, 8);
                p += __shfl_xor_sync(kFull, p, 16);
                if (c == 0 && n < kMaxNodes) {
                    float x = -INFINITY;
                    if (live) {
                        x = __fdiv_rn(p, sqrt_e);
                        if (isnan(x)) atomicOr(&misc[M_STATUS], RL4CO_ST_NAN_LOGITS);
                        if (a.tanh_clipping > 0.f) x = tanhf(x) * a.tanh_clipping;
                        x = __fdiv_rn(x, inv_temp_div);
                    }
                    sx[n] = x;
                }
            }
            __syncthreads();

            // ---- masked log-softmax + action selection + env transition: warp 0
            if (warp == 0) {
                float v[4], m = -INFINITY;
#pragma unroll
                for (int k = 0; k < 4; ++k) { v[k] = sx[lane * 4 + k]; m = fmaxf(m, v[k]); }
                m = warp_max(m);
                float e[4], lsum = 0.f;
#pragma unroll
                for (int k = 0; k < 4; ++k) { e[k] = (v[k] == -INFINITY) ? 0.f : expf(v[k] - m); lsum += e[k]; }
                const float total = warp_sum(lsum);
                const float lse = logf(total);
                int act;
                if (a.mode == RL4CO_MODE_EVALUATE) {
                    int64_t g = a.actions_in[(size_t)r * T + t];
                    if (g < 0 || g >= N) { if (lane == 0) atomicOr(&misc[M_STATUS], RL4CO_ST_ACTION_RANGE); g = g < 0 ? 0 : N - 1; }
                    act = (int)g;
                } else if (a.mode == RL4CO_MODE_GREEDY) {
                    // argmax, lowest index on ties (torch.argmax on CPU / first max)
                    float bv = v[0]; int bi = lane * 4;
#pragma unroll
                    for (int k = 1; k < 4; ++k) if (v[k] > bv) { bv = v[k]; bi = lane * 4 + k; }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        float ov = __shfl_xor_sync(kFull, bv, o);
                        int oi = __shfl_xor_sync(kFull, bi, o);
                        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
                    }
                    act = bi;
                } else {
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
                const float xa = sx[act];
                const float lpa = (xa - m) - lse;
                if (a.logp_all) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        int nn = lane * 4 + k;
                        if (nn < N) a.logp_all[((size_t)r * T + t) * N + nn] = (v[k] - m) - lse;
                    }
                }
                if (lane == 0) {
                    int st = 0;
                    if (!feas[act]) st |= RL4CO_ST_INFEASIBLE_ACTION;
                    if (!(lpa > -1000.f)) st |= RL4CO_ST_NEG_INF_LOGP;
                    if (st) atomicOr(&misc[M_STATUS], st);
                    s_act[t] = act; s_lp[t] = lpa; ll += lpa;
                    misc[M_NACT] = t + 1;
                }
                transition(act);
            }
            __syncthreads();
        }

        // ================================================== epilogue =====================
        const int nact = misc[M_NACT];
        float part = 0.f;
        if (ENV == RL4CO_ENV_TSP) {              // tsp/env.py:157-163
            for (int t = tid; t < nact; t += kThreads) {
                float2 p = s_locs[s_act[t]], q = s_locs[s_act[t + 1 == nact ? 0 : t + 1]];
                part += leg(p.x, p.y, q.x, q.y);
            }
        } else if (ENV == RL4CO_ENV_CVRP) {      // cvrp/env.py:151-160 (depot at both ends)
            for (int t = tid; t <= nact; t += kThreads) {
                float2 p = t == 0 ? s_locs[0] : s_locs[s_act[t - 1]];
                float2 q = t == nact ? s_locs[0] : s_locs[s_act[t]];
                part += leg(p.x, p.y, q.x, q.y);
            }
        } else {                                 // slap/env.py:131-143 (assignment[t] = s_act[t])
            const int64_t *pl = a.picklist + b * (int64_t)a.O * a.K;
            for (int o = tid; o < a.O; o += kThreads)
                part += slap_order_len(s_locs, s_act, pl + (int64_t)o * a.K, a.K);
        }
        part = warp_sum(part);
        if (lane == 0) spart[warp] = part;
        __syncthreads();
        if (tid == 0) {
            float len = 0.f;
            for (int g = 0; g < kWarps; ++g) len += spart[g];
            a.reward[r] = -len;
            a.ll[r] = ll;
            if (a.steps) a.steps[r] = nact;
            if (misc[M_STATUS]) raise_status(a.status, misc[M_STATUS]);
        }
        for (int t = tid; t < T; t += kThreads) {    // pad with depot / log-prob 0 after done
            a.actions[(size_t)r * T + t] = t < nact ? (int64_t)s_act[t] : 0;
            if (a.logp) a.logp[(size_t)r * T + t] = t < nact ? s_lp[t] : 0.f;
        }
        if (a.logp_all)
            for (size_t k = (size_t)nact * N + tid; k < (size_t)T * N; k += kThreads)
                a.logp_all[(size_t)r * T * N + k] = 0.f;
        if (ENV == RL4CO_ENV_SLAP && a.assignment)
            for (int t = tid; t < a.P; t += kThreads) a.assignment[(size_t)r * a.P + t] = t < nact ? s_act[t] : -1;
        __syncthreads();
    }
}

}  // namespace rl4co

using namespace rl4co;

template <int ENV>
static int launch_rollout(const rl4co_am_rollout_args &a, int n_smem_tab, size_t smem,
                          cudaStream_t stream) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (sm_count <= 0) sm_count = 148;
    }
    cudaError_t e = cudaFuncSetAttribute(am_rollout_kernel<ENV>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        set_error("am_rollout: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e));
        return 2;
    }
    const int S = a.S > 1 ? a.S : 1;
    const int64_t R = a.B * (int64_t)S;
    // persistent grid: one CTA per SM (the tables fill the SM's shared memory), contiguous
    // chunks of work items so that the starts of one instance reuse the staged tables
    int64_t grid = R < sm_count ? R : sm_count;
    int64_t chunk = (R + grid - 1) / grid;
    grid = (R + chunk - 1) / chunk;
    am_rollout_kernel<ENV><<<(unsigned)grid, kThreads, smem, stream>>>(a, n_smem_tab, R, chunk);
    RL4CO_CHECK_LAUNCH("am_rollout");
    return 0;
}

static int g_rollout_impl = 0;
extern "C" int rl4co_set_rollout_impl(int impl) {
    RL4CO_REQUIRE(impl >= 0 && impl <= 2, "set_rollout_impl: 0 = automatic, 1 = shared-memory kernel only, 2 = no TSP-greedy specialisation");
    g_rollout_impl = impl;
    return 0;
}

extern "C" int rl4co_am_rollout(const rl4co_am_rollout_args *args, rl4co_stream_t stream) {
    RL4CO_REQUIRE(args != nullptr, "am_rollout: null args");
    const rl4co_am_rollout_args &a = *args;
    if (a.B == 0) return 0;
    RL4CO_REQUIRE(a.env >= RL4CO_ENV_TSP && a.env <= RL4CO_ENV_SDVRP, "am_rollout: unknown env %d", a.env);
    RL4CO_REQUIRE(a.mode >= RL4CO_MODE_GREEDY && a.mode <= RL4CO_MODE_EVALUATE, "am_rollout: unknown mode %d", a.mode);
    RL4CO_REQUIRE(a.B >= 0 && a.N >= 2 && a.T >= 1, "am_rollout: bad sizes B=%lld N=%d T=%d", (long long)a.B, a.N, a.T);
    RL4CO_REQUIRE(a.N <= kMaxNodes, "am_rollout: N=%d exceeds the %d node slots of this kernel", a.N, kMaxNodes);
    RL4CO_REQUIRE(a.tables && a.locs && a.actions && a.ll && a.reward, "am_rollout: null tensor");
    RL4CO_REQUIRE(a.temperature > 0.f, "am_rollout: temperature must be > 0");
    RL4CO_REQUIRE(a.mode != RL4CO_MODE_EVALUATE || a.actions_in, "am_rollout: evaluate needs actions_in");
    // tsp: table 4 may be left out when the kernel can form its one row itself (h and w_first given)
    RL4CO_REQUIRE((a.h == nullptr) == (a.w_first == nullptr), "am_rollout: h and w_first come together");
    RL4CO_REQUIRE(a.w_first == nullptr || a.env == RL4CO_ENV_TSP, "am_rollout: h / w_first are tsp arguments");
    RL4CO_REQUIRE(a.w_first == nullptr || ((((uintptr_t)a.h) | ((uintptr_t)a.w_first)) & 15) == 0,
                  "am_rollout: h / w_first must be 16-byte aligned");
    const int need_tab = a.env == RL4CO_ENV_TSP ? (a.w_first ? 4 : 5) : (a.env == RL4CO_ENV_SLAP ? 3 : 4);
    RL4CO_REQUIRE(a.n_tables >= need_tab, "am_rollout: env needs %d tables, got %d", need_tab, a.n_tables);
    if (a.env == RL4CO_ENV_TSP) RL4CO_REQUIRE(a.q0, "am_rollout: tsp needs q0");
    if (a.env == RL4CO_ENV_CVRP)
        RL4CO_REQUIRE(a.wcap && a.demand && a.vehicle_capacity, "am_rollout: cvrp needs wcap/demand/capacity");
    if (a.env == RL4CO_ENV_SLAP)
        RL4CO_REQUIRE(a.picklist && a.P >= 1 && a.O >= 0 && a.K >= 1 && a.T >= a.P, "am_rollout: slap needs picklist and T>=P");
    if (a.env == RL4CO_ENV_OP) {
        RL4CO_REQUIRE(a.wcap && a.prize && a.max_length, "am_rollout: op needs wcap/prize/max_length");
        // multistart: where the reference resamples the start nodes (utils/ops.py:166-176) the caller passes them in start_nodes
        RL4CO_REQUIRE(!a.multistart || a.num_loc >= 1, "am_rollout: bad multistart setup");
        RL4CO_REQUIRE(rollout_rf_supports(a), "am_rollout: op runs on the register-file kernel only (N <= 128, no logp_all)");
        return launch_rollout_rf(a, (cudaStream_t)stream);
    }
    if (a.env == RL4CO_ENV_SDVRP) {
        RL4CO_REQUIRE(a.wcap && a.demand && a.vehicle_capacity && a.wdyn, "am_rollout: sdvrp needs wcap/demand/capacity/wdyn");
        RL4CO_REQUIRE(!a.multistart || a.num_loc >= 1, "am_rollout: bad multistart setup");
        RL4CO_REQUIRE(rollout_rf_supports(a), "am_rollout: sdvrp runs on the register-file kernel only (N <= 128, no logp_all)");
        return launch_rollout_rf(a, (cudaStream_t)stream);
    }
    if (a.env == RL4CO_ENV_PCTSP) {
        RL4CO_REQUIRE(a.wcap && a.prize && a.penalty && a.prize_required, "am_rollout: pctsp needs wcap/prize/penalty/prize_required");
        RL4CO_REQUIRE(!a.multistart || a.num_loc >= 1, "am_rollout: bad multistart setup");
        RL4CO_REQUIRE(rollout_rf_supports(a), "am_rollout: pctsp runs on the register-file kernel only (N <= 128, no logp_all)");
        return launch_rollout_rf(a, (cudaStream_t)stream);
    }
    if (a.multistart) RL4CO_REQUIRE(a.num_loc >= 1 && a.env != RL4CO_ENV_SLAP, "am_rollout: bad multistart setup");
    RL4CO_REQUIRE(a.start_nodes == nullptr || a.multistart, "am_rollout: start_nodes without multistart");

    // multistart (POMO) tsp / cvrp: tensor-core kernel, starts = TMEM lanes (am_rollout_ms_tc.cu)
    if (g_rollout_impl == 0 && rollout_ms_tc_supports(a)) return launch_rollout_ms_tc(a, (cudaStream_t)stream);
    // TSP greedy single start: register-file kernel with node compaction (am_rollout_tsp.cu)
    if (g_rollout_impl == 0 && rollout_tsp_greedy_supports(a)) return launch_rollout_tsp_greedy(a, (cudaStream_t)stream);
    // slap: the query never changes, so the glimpse is maintained incrementally (am_rollout_slap.cu)
    if (g_rollout_impl == 0 && rollout_slap_supports(a)) return launch_rollout_slap(a, (cudaStream_t)stream);
    // single-start rollouts: register-file resident kernel (am_rollout_rf.cu)
    if (g_rollout_impl != 1 && rollout_rf_supports(a)) return launch_rollout_rf(a, (cudaStream_t)stream);

    int n_smem_tab = a.env == RL4CO_ENV_SLAP ? 3 : 4;
    const size_t limit = 227 * 1024;
    Layout L = make_layout(a.N, a.T, n_smem_tab);
    if (L.total > limit && n_smem_tab == 4) {    // keep the context table in global (L2) instead
        n_smem_tab = 3;
        L = make_layout(a.N, a.T, n_smem_tab);
    }
    RL4CO_REQUIRE(L.total <= limit, "am_rollout: N=%d T=%d needs %zu B shared memory (> %zu)", a.N, a.T, L.total, limit);
    cudaStream_t s = (cudaStream_t)stream;
    switch (a.env) {
        case RL4CO_ENV_TSP: return launch_rollout<RL4CO_ENV_TSP>(a, n_smem_tab, L.total, s);
        case RL4CO_ENV_CVRP: return launch_rollout<RL4CO_ENV_CVRP>(a, n_smem_tab, L.total, s);
        default: return launch_rollout<RL4CO_ENV_SLAP>(a, n_smem_tab, L.total, s);
    }
}
