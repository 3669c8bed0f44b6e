// K1 / K2: field-wise embedding gather fused with the FM first/second-order terms.
//
// One warp per sample.  A row of k floats is fetched by k/4 lanes with one 128-bit load each, so a
// warp-wide load instruction brings in 32/(k/4) table rows at once and the same registers feed
//   (1) the FM accumulators  sum_f e  and  sum_f e^2,
//   (2) the deep-tower input tile X0 (written as one contiguous, fully coalesced run per warp store),
//   (3) the first-order output.
// Every table row is therefore read from HBM exactly once per lookup (the reference gathers
// feats_emb twice and feats once: models/deepfm.py:78,85,98).
#include "ctr_common.cuh"
#include <cstdlib>
#include <type_traits>

namespace {

struct f4 {
    float x, y, z, w;
};

__device__ __forceinline__ float4 f4_zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ float4 f4_nan() {
    const float n = __int_as_float(0x7fc00000);
    return make_float4(n, n, n, n);
}
__device__ __forceinline__ void f4_acc(float4& s, float4& q, const float4 e) {
    s.x += e.x; s.y += e.y; s.z += e.z; s.w += e.w;
    q.x = fmaf(e.x, e.x, q.x); q.y = fmaf(e.y, e.y, q.y); q.z = fmaf(e.z, e.z, q.z); q.w = fmaf(e.w, e.w, q.w);
}
__device__ __forceinline__ float4 f4_shfl_xor_add(float4 v, int o) {
    v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
    v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
    v.z += __shfl_xor_sync(0xffffffffu, v.z, o);
    v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
    return v;
}

template <int K, bool USE_FM>
__global__ void __launch_bounds__(256)
fm_gather_fwd_kernel(const ctr_embed_cfg cfg, const float* __restrict__ table, const void* __restrict__ ids,
                     const int64_t id_offset, const float* __restrict__ cont, const float* __restrict__ vec,
                     const float* __restrict__ cont_rows, const float* __restrict__ pooled_first,
                     float* __restrict__ x0, float* __restrict__ fm, float* __restrict__ sum_e) {
    constexpr int LPR = K / 4;       // lanes per row
    constexpr int RPW = 32 / LPR;    // rows per warp-wide load
    constexpr int UNR = 4;           // independent row loads in flight per lane
    const int lane = threadIdx.x & 31;
    const int64_t b = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= cfg.B) return;
    const int r = lane / LPR, q = lane % LPR;
    const int Fc = cfg.Fc, Fm = cfg.Fm, C = cfg.C, V = cfg.V;
    const int stride = cfg.table_stride;
    const int64_t idrow = b * (int64_t) cfg.ids_ld;
    float* __restrict__ x0row = x0 + b * (int64_t) cfg.x0_ld;
    float* __restrict__ fmrow = USE_FM ? fm + b * (int64_t) cfg.fm_ld : nullptr;
    const bool has_first = cfg.has_first != 0;
    const bool zero0 = cfg.zero_row0 != 0;
    const bool rnd = cfg.round_tf32 != 0;   // TF32 tower: pre-round the tile the tensor cores will read

    float4 as = f4_zero(), aq = f4_zero();

    // ---- single-valued fields: gather ------------------------------------------------
    for (int f0 = 0; f0 < Fc; f0 += RPW * UNR) {
        int64_t rid[UNR];
        bool ok[UNR], live[UNR], bad[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            int f = f0 + u * RPW + r;
            ok[u] = f < Fc;
            rid[u] = ok[u] ? ctr_load_id(ids, cfg.id_bytes, idrow + f) + id_offset : 0;
            bad[u] = ok[u] && cfg.table_rows > 0 && (uint64_t) rid[u] >= (uint64_t) cfg.table_rows;
            if (bad[u]) ctr_flag_bad_id(cfg.status, f);
            live[u] = ok[u] && !bad[u] && !(zero0 && rid[u] == 0);
        }
        float4 e[UNR];
        float w[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            const float* row = table + rid[u] * (int64_t) stride;
            e[u] = live[u] ? __ldg(reinterpret_cast<const float4*>(row + q * 4)) : f4_zero();
            w[u] = (USE_FM && live[u] && has_first && q == 0) ? __ldg(row + K) : 0.f;
            if (bad[u]) e[u] = f4_nan();      // tf.nn.embedding_lookup raises here: the sample's logit / loss / score turn NaN
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            if (ok[u]) {
                int f = f0 + u * RPW + r;
                *reinterpret_cast<float4*>(x0row + f * K + q * 4) =
                    rnd ? make_float4(ctr_round_tf32(e[u].x), ctr_round_tf32(e[u].y), ctr_round_tf32(e[u].z), ctr_round_tf32(e[u].w)) : e[u];
                if (USE_FM) {
                    f4_acc(as, aq, e[u]);
                    if (q == 0) fmrow[f] = w[u];
                }
            }
        }
    }

    // ---- continuous / vector values -> X0 ----------------------------------------------
    const int cont_col = (Fc + Fm) * K;
    for (int c = lane; c < C; c += 32) {
        const float x = cont[b * (int64_t) C + c];
        x0row[cont_col + c] = rnd ? ctr_round_tf32(x) : x;
    }
    for (int j = lane; j < V; j += 32) {
        const float x = vec[b * (int64_t) V + j];
        x0row[cont_col + C + j] = rnd ? ctr_round_tf32(x) : x;
    }
    // pad columns: a 1.0 at column D0 lets the layer-0 weight-gradient GEMM produce the bias gradient as an extra
    // row (ctr_mlp_bwd_weight x_ones_col); the forward GEMM reads exactly D0 columns and never sees it
    for (int j = cont_col + C + V + lane; j < cfg.x0_ld; j += 32) x0row[j] = (j == cont_col + C + V) ? 1.f : 0.f;

    if (!USE_FM) return;

    // ---- pooled multi-valued fields enter the FM as one field each (models/deepfm_multi.py:165-167)
    for (int j0 = 0; j0 < Fm; j0 += RPW) {
        int j = j0 + r;
        if (j < Fm) {
            float4 e = *reinterpret_cast<const float4*>(x0row + (Fc + j) * K + q * 4);
            f4_acc(as, aq, e);
            if (q == 0) fmrow[Fc + j] = pooled_first ? pooled_first[b * (int64_t) Fm + j] : 0.f;
        }
    }

    // ---- continuous fields: e = x * theta_c  (models/deepfm.py:85-87) ------------------------
    if (cfg.fm_cont) {
        for (int c0 = 0; c0 < C; c0 += RPW) {
            int c = c0 + r;
            if (c < C) {
                float x = cont[b * (int64_t) C + c];
                const float* row = cont_rows + c * (int64_t) stride;
                float4 t = __ldg(reinterpret_cast<const float4*>(row + q * 4));
                float4 e = make_float4(t.x * x, t.y * x, t.z * x, t.w * x);
                f4_acc(as, aq, e);
                if (q == 0) fmrow[Fc + Fm + c] = (has_first ? __ldg(row + K) : 0.f) * x;
            }
        }
    } else {
        for (int c = lane; c < C; c += 32) fmrow[Fc + Fm + c] = 0.f;
    }

    // ---- reduce the row slots, finish 0.5 * ((sum e)^2 - sum e^2)  (models/deepfm.py:88-92)
#pragma unroll
    for (int o = LPR; o < 32; o <<= 1) {
        as = f4_shfl_xor_add(as, o);
        aq = f4_shfl_xor_add(aq, o);
    }
    if (r == 0) {
        float* sec = fmrow + Fc + Fm + C + q * 4;
        sec[0] = 0.5f * (as.x * as.x - aq.x);
        sec[1] = 0.5f * (as.y * as.y - aq.y);
        sec[2] = 0.5f * (as.z * as.z - aq.z);
        sec[3] = 0.5f * (as.w * as.w - aq.w);
        *reinterpret_cast<float4*>(sum_e + b * (int64_t) K + q * 4) = as;
    }
}

// K1, fast path (no pooled fields, no vector block, C <= 2 warp-loads of rows): the same lane mapping as above,
// restated so that a sample costs ~200 warp instructions instead of ~660 -- the general kernel is issue-bound, not
// HBM-bound (profiles/r01a: sm__inst_executed 55 % of peak at 45 % occupancy, DRAM 24 %).
//   * warps are persistent over a grid-stride sequence of samples, so everything that is per-model rather than
//     per-sample (flags, strides, the C continuous-field rows, which stay in registers) is set up once;
//   * the ids and continuous values of the NEXT sample are fetched (one coalesced load each) before the current
//     sample's rows are requested, which takes the id -> row dependent-load chain off the critical path;
//   * lane l owns id l: it fetches that row's first-order weight itself, so the first-order run is one coalesced
//     store; the row-gather lanes receive their ids by shuffle;
//   * the (sum e, sum e^2) reduction over the row slots is a halving exchange (7 shuffles instead of 24).
template <int K, bool USE_FM, bool ID64>
__global__ void __launch_bounds__(256, 4)
fm_gather_fast_kernel(const ctr_embed_cfg cfg, const float* __restrict__ table, const void* __restrict__ ids,
                      const int64_t id_offset, const float* __restrict__ cont, const float* __restrict__ cont_rows,
                      float* __restrict__ x0, float* __restrict__ fm, float* __restrict__ sum_e) {
    constexpr int LPR = K / 4;                                   // lanes per row
    constexpr int RPW = 32 / LPR;                                // rows per warp-wide load
    constexpr int UNR = LPR;                                     // warp-wide loads per 32 fields
    constexpr int CS = 2;                                        // continuous-row slots per lane
    constexpr int H = RPW >= 8 ? 3 : 2;                          // halving levels (offsets 16, 8[, 4])
    constexpr int NV = 8 >> H;                                   // values left per lane after them
    constexpr int BF_MASK = LPR < 4 ? (3 & ~(LPR - 1)) : 0;      // lane bits reduced by plain butterflies
    typedef typename std::conditional<ID64, int64_t, int32_t>::type id_t;
    const int lane = threadIdx.x & 31;
    const int r = lane / LPR, q = lane % LPR;
    const int Fc = cfg.Fc, C = cfg.C, B = cfg.B;
    const int stride = cfg.table_stride;
    const bool has_first = cfg.has_first != 0, zero0 = cfg.zero_row0 != 0, rnd = cfg.round_tf32 != 0;
    const bool fm_cont = USE_FM && cfg.fm_cont != 0;
    const int cont_col = Fc * K;
    const int tail = cfg.x0_ld - cont_col;      // continuous values + pad columns (<= 32, checked by the launcher)
    const int nw = (int) ((gridDim.x * blockDim.x) >> 5);
    int b = (int) ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (b >= B) return;

    float4 crow[CS];
    float cw = 0.f;
#pragma unroll
    for (int s = 0; s < CS; ++s) {
        const int c = s * RPW + r;
        crow[s] = (fm_cont && c < C) ? __ldg(reinterpret_cast<const float4*>(cont_rows + c * (int64_t) stride + q * 4)) : f4_zero();
    }
    if (fm_cont && has_first && lane < C) cw = __ldg(cont_rows + lane * (int64_t) stride + K);

    const id_t* __restrict__ idp = reinterpret_cast<const id_t*>(ids);
    const char* __restrict__ tbytes = reinterpret_cast<const char*>(table);
    const uint32_t stride_b = (uint32_t) stride * 4u;
    id_t my_id = lane < Fc ? idp[b * (int64_t) cfg.ids_ld + lane] : (id_t) 0;
    float my_x = lane < C ? cont[b * (int64_t) C + lane] : 0.f;

    for (; b < B; b += nw) {
        const int bn = b + nw;
        id_t nx_id = 0;
        float nx_x = 0.f;
        if (bn < B) {
            if (lane < Fc) nx_id = idp[bn * (int64_t) cfg.ids_ld + lane];
            if (lane < C) nx_x = cont[bn * (int64_t) C + lane];
        }
        float* __restrict__ x0row = x0 + b * (int64_t) cfg.x0_ld;
        float* __restrict__ fmrow = USE_FM ? fm + b * (int64_t) cfg.fm_ld : nullptr;
        float4 as = f4_zero(), aq = f4_zero();

        for (int f0 = 0; f0 < Fc; f0 += 32) {
            id_t id_l = my_id;
            if (f0 > 0) id_l = f0 + lane < Fc ? idp[b * (int64_t) cfg.ids_ld + f0 + lane] : (id_t) 0;
            // row index as u32 (tables have < 2^32 rows); lanes past Fc carry id 0 -> a valid row that is never used
            uint32_t row_l = (uint32_t) ((int64_t) id_l + id_offset);
            // an id outside the table (tf.nn.embedding_lookup raises): no memory is read for it, the row reads as NaN
            const bool bad_l = cfg.table_rows > 0 && (uint64_t) ((int64_t) id_l + id_offset) >= (uint64_t) cfg.table_rows;
            const uint32_t badmask = __ballot_sync(0xffffffffu, bad_l);
            if (bad_l) {
                ctr_flag_bad_id(cfg.status, f0 + lane);
                row_l = 0;
            }
            float4 e[UNR];
            uint32_t row[UNR];
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                row[u] = __shfl_sync(0xffffffffu, row_l, u * RPW + r);
                // L1-allocating loads on purpose: the hot rows of a skewed id distribution are served from L1; sent
                // to L2 every time (ld.global.nc.L1::no_allocate) they queue on a few L2 slices and the kernel is 1.5x slower
                e[u] = __ldg(reinterpret_cast<const float4*>(tbytes + (uint64_t) row[u] * stride_b) + q);
            }
            if (USE_FM) {      // first-order run: lane l owns field f0 + l
                float w = 0.f;
                if (has_first) w = __ldg(reinterpret_cast<const float*>(tbytes + (uint64_t) row_l * stride_b) + K);
                if (zero0 && row_l == 0) w = 0.f;
                if (f0 + lane < Fc) fmrow[f0 + lane] = w;
            }
            if (zero0) {
#pragma unroll
                for (int u = 0; u < UNR; ++u)
                    if (row[u] == 0) e[u] = f4_zero();
            }
            if (badmask) {             // warp-uniform, error path only
#pragma unroll
                for (int u = 0; u < UNR; ++u)
                    if ((badmask >> (u * RPW + r)) & 1u) e[u] = f4_nan();
            }
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                if (f0 + u * RPW + r < Fc) {
                    float4 o = e[u];
                    if (rnd) o = make_float4(ctr_round_tf32_fast(o.x), ctr_round_tf32_fast(o.y), ctr_round_tf32_fast(o.z), ctr_round_tf32_fast(o.w));
                    *reinterpret_cast<float4*>(x0row + (f0 + u * RPW) * K + lane * 4) = o;
                    if (USE_FM) f4_acc(as, aq, e[u]);
                }
            }
        }
        // continuous values, then the pad columns: the 1.0 at column D0 is the bias row of the layer-0 weight-gradient GEMM
        if (lane < tail) x0row[cont_col + lane] = lane < C ? (rnd ? ctr_round_tf32_fast(my_x) : my_x) : (lane == C ? 1.f : 0.f);

        if (USE_FM) {
            if (lane < C) fmrow[Fc + lane] = cw * my_x;          // cw is 0 without fm_cont / first-order weights
            if (fm_cont) {
#pragma unroll
                for (int s = 0; s < CS; ++s) {
                    const float x = __shfl_sync(0xffffffffu, my_x, (s * RPW + r) & 31);     // slots c >= C hold zero rows
                    f4_acc(as, aq, make_float4(crow[s].x * x, crow[s].y * x, crow[s].z * x, crow[s].w * x));
                }
            }
            float v[8] = {as.x, as.y, as.z, as.w, aq.x, aq.y, aq.z, aq.w};
#pragma unroll
            for (int j = 0; j < H; ++j) {        // halving exchange: each level keeps half of the values, summed over the pair
                const int o = 16 >> j, n = 4 >> j;
                const bool hi = (lane & o) != 0;
#pragma unroll
                for (int i = 0; i < n; ++i) {
                    const float keep = hi ? v[i + n] : v[i], send = hi ? v[i] : v[i + n];
                    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
                }
            }
#pragma unroll
            for (int o = 16 >> H; o >= LPR; o >>= 1) {
#pragma unroll
                for (int i = 0; i < NV; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
            }
            // lanes with bit 16 clear hold sum e, their partners sum e^2, of coordinate q*4 + comp
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const float sq = __shfl_xor_sync(0xffffffffu, v[i], 16);
                if ((lane & (16 | BF_MASK)) == 0) {
                    const int comp = (((lane & 8) ? 2 : 0) | ((H >= 3 && (lane & 4)) ? 1 : 0)) + i;
                    const int j = q * 4 + comp;
                    fmrow[Fc + C + j] = 0.5f * (v[i] * v[i] - sq);      // models/deepfm.py:88-92
                    sum_e[b * (int64_t) K + j] = v[i];
                }
            }
        }
        my_id = nx_id;
        my_x = nx_x;
    }
}

// ---- async-copy plumbing of the staged K1 ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t) __cvta_generic_to_shared(p); }
// LDGSTS: global -> shared without a register round trip; src_bytes < size zero-fills the rest (0 = a row of zeros
// without touching memory: padding id 0 of the row-0-zero models, ids outside the table)
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// K1, staged path (same shapes as the fast path above, Fc <= 32).  The register-staged kernel above is ISSUE-bound
// (ncu, profiles/r01h: 403 warp instructions per sample, issue slots 60 % busy, DRAM 39 %) and its bytes in flight are
// capped by the register file.  Here a warp owns a ring of NS gathered-row images in shared memory:
//   fill     the 16-byte chunks of a sample's rows go global -> shared with cp.async (LDGSTS): no destination
//            registers, zero-fill for padding ids; the first-order weights ride along as 4-byte copies;
//   consume  LAG samples later each lane reads its chunks back (conflict-free 128-bit LDS), accumulates the FM sums and
//            stores the (TF32-rounded) chunk to X0 with an immediate offset from ONE per-sample base pointer.
// The loads of NS-1 samples (1.7 KB each) are in flight per warp whatever the register budget.
// (A TMA bulk store of the finished row image, cp.async.bulk shared -> global, was tried first: its bulk groups count on
// the same scoreboard as the LDGSTS groups -- both wait with DEPBAR.LE SB0 -- so every fill wait also waited for the
// previous row's store; and it needs the rounded values written back to shared memory plus a proxy fence per row.)
template <int K, bool USE_FM, bool ID64, int NS>
__global__ void __launch_bounds__(256)
fm_gather_stage_kernel(const ctr_embed_cfg cfg, const float* __restrict__ table, const void* __restrict__ ids,
                       const int64_t id_offset, const float* __restrict__ cont, const float* __restrict__ cont_rows,
                       float* __restrict__ x0, float* __restrict__ fm, float* __restrict__ sum_e) {
    constexpr int LPR = K / 4;                                   // lanes (16-byte chunks) per row
    constexpr int RPW = 32 / LPR;                                // rows per warp-wide chunk sweep
    constexpr int CS = 2;                                        // continuous-row slots per lane
    constexpr int LAG = NS - 1;                                  // samples between fill and consume
    constexpr int H = RPW >= 8 ? 3 : 2;
    constexpr int NV = 8 >> H;
    constexpr int BF_MASK = LPR < 4 ? (3 & ~(LPR - 1)) : 0;
    constexpr int XTRA = 36;                                     // per slot behind the rows: 32 first-order words, 1 fault mask
    typedef typename std::conditional<ID64, int64_t, int32_t>::type id_t;
    extern __shared__ __align__(16) float k1_smem[];
    const int lane = threadIdx.x & 31;
    const int r = lane / LPR, q = lane % LPR;
    const int Fc = cfg.Fc, C = cfg.C, B = cfg.B;
    const bool has_first = cfg.has_first != 0, zero0 = cfg.zero_row0 != 0, rnd = cfg.round_tf32 != 0;
    const bool fm_cont = USE_FM && cfg.fm_cont != 0;
    const int cont_col = Fc * K;
    const int tail = cfg.x0_ld - cont_col;      // continuous values + pad columns (<= 32, checked by the launcher)
    const int slot_f = cont_col + XTRA;
    float* const wbase = k1_smem + (size_t) ((threadIdx.x >> 5) * NS) * slot_f;
    const int nw = (int) ((gridDim.x * blockDim.x) >> 5);
    const int b0 = (int) ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (b0 >= B) return;
    const int n_local = (B - b0 + nw - 1) / nw;

    const int stride = cfg.table_stride;
    float4 crow[CS];
    float cw = 0.f;
#pragma unroll
    for (int s = 0; s < CS; ++s) {
        const int c = s * RPW + r;
        crow[s] = (fm_cont && c < C) ? __ldg(reinterpret_cast<const float4*>(cont_rows + c * (int64_t) stride + q * 4)) : f4_zero();
    }
    if (fm_cont && has_first && lane < C) cw = __ldg(cont_rows + lane * (int64_t) stride + K);

    const id_t* __restrict__ idp = reinterpret_cast<const id_t*>(ids);
    const char* __restrict__ tbytes = reinterpret_cast<const char*>(table);
    const uint32_t stride_b = (uint32_t) stride * 4u;
    const uint32_t s_base = smem_u32(wbase);
    id_t nx_id = lane < Fc ? idp[b0 * (int64_t) cfg.ids_ld + lane] : (id_t) 0;
    float nx_x = lane < C ? cont[b0 * (int64_t) C + lane] : 0.f;

    for (int t = 0; t < n_local + LAG; ++t) {
        // ---------------------------------------------------------------- fill: sample t -> ring slot t % NS
        // (the slot was last read by this warp's consume of sample t - NS, earlier in program order)
        if (t < n_local) {
            const int b = b0 + t * nw;
            const id_t id_l = nx_id;
            if (t + 1 < n_local && lane < Fc) nx_id = idp[(b + nw) * (int64_t) cfg.ids_ld + lane];
            const int64_t rix = (int64_t) id_l + id_offset;
            const bool bad_l = lane < Fc && cfg.table_rows > 0 && (uint64_t) rix >= (uint64_t) cfg.table_rows;
            const uint32_t row_l = bad_l ? 0u : (uint32_t) rix;
            const uint32_t zmask = __ballot_sync(0xffffffffu, bad_l || lane >= Fc || (zero0 && row_l == 0u));
            const uint32_t badmask = __ballot_sync(0xffffffffu, bad_l);
            const uint32_t s_slot = s_base + (uint32_t) ((t % NS) * slot_f) * 4u;
#pragma unroll
            for (int j = 0; j < LPR; ++j) {                              // Fc <= 32: at most LPR sweeps of 32 chunks
                const int f = r + j * RPW;                               // field of chunk lane + 32 j
                const uint32_t rw = __shfl_sync(0xffffffffu, row_l, f & 31);
                if (f < Fc)
                    cp_async16(s_slot + (uint32_t) (lane + 32 * j) * 16u, tbytes + (uint64_t) rw * stride_b + q * 16,
                               ((zmask >> f) & 1u) ? 0u : 16u);
            }
            if (USE_FM && lane < Fc)             // first-order weight of field `lane` (0 without the column / for padding)
                cp_async4(s_slot + (uint32_t) (cont_col + lane) * 4u, tbytes + (uint64_t) row_l * stride_b + K * 4,
                          (has_first && !((zmask >> lane) & 1u)) ? 4u : 0u);
            if (lane == 0) wbase[(t % NS) * slot_f + cont_col + 32] = __uint_as_float(badmask);
            if (bad_l) ctr_flag_bad_id(cfg.status, lane);
        }
        cp_async_commit();                       // one group per iteration (empty in the drain) keeps the wait count fixed
        // ---------------------------------------------------------------- consume: sample t - LAG
        const int tc = t - LAG;
        if (tc < 0) continue;                    // (warp-uniform)
        cp_async_wait<LAG>();
        __syncwarp();
        const int b = b0 + tc * nw;
        const float* const slot = wbase + (tc % NS) * slot_f;
        float* __restrict__ x0row = x0 + b * (int64_t) cfg.x0_ld;
        const float my_x = nx_x;
        if (tc + 1 < n_local && lane < C) nx_x = cont[(b + nw) * (int64_t) C + lane];
        const uint32_t badmask = __float_as_uint(slot[cont_col + 32]);
        float4 as = f4_zero(), aq = f4_zero();
#pragma unroll
        for (int j = 0; j < LPR; ++j) {
            const int f = r + j * RPW;
            if (f < Fc) {
                float4 e = reinterpret_cast<const float4*>(slot)[lane + 32 * j];
                if (badmask && ((badmask >> f) & 1u)) e = f4_nan();      // an id outside the table: tf.nn.embedding_lookup raises
                float4 o = e;
                if (rnd) o = make_float4(ctr_round_tf32_fast(e.x), ctr_round_tf32_fast(e.y), ctr_round_tf32_fast(e.z), ctr_round_tf32_fast(e.w));
                reinterpret_cast<float4*>(x0row)[lane + 32 * j] = o;
                if (USE_FM) f4_acc(as, aq, e);
            }
        }
        // continuous values, then the pad columns: the 1.0 at column D0 is the bias row of the layer-0 weight-gradient GEMM
        if (lane < tail) x0row[cont_col + lane] = lane < C ? (rnd ? ctr_round_tf32_fast(my_x) : my_x) : (lane == C ? 1.f : 0.f);
        if (USE_FM) {
            float* __restrict__ fmrow = fm + b * (int64_t) cfg.fm_ld;
            if (lane < Fc) {
                float w = slot[cont_col + lane];
                if (badmask && ((badmask >> lane) & 1u)) w = __int_as_float(0x7fc00000);
                fmrow[lane] = w;
            }
            if (lane < C) fmrow[Fc + lane] = cw * my_x;          // cw is 0 without fm_cont / first-order weights
            if (fm_cont) {
#pragma unroll
                for (int s = 0; s < CS; ++s) {
                    const float x = __shfl_sync(0xffffffffu, my_x, (s * RPW + r) & 31);     // slots c >= C hold zero rows
                    f4_acc(as, aq, make_float4(crow[s].x * x, crow[s].y * x, crow[s].z * x, crow[s].w * x));
                }
            }
            float v[8] = {as.x, as.y, as.z, as.w, aq.x, aq.y, aq.z, aq.w};
#pragma unroll
            for (int j = 0; j < H; ++j) {        // halving exchange: each level keeps half of the values, summed over the pair
                const int o = 16 >> j, n = 4 >> j;
                const bool hi = (lane & o) != 0;
#pragma unroll
                for (int i = 0; i < n; ++i) {
                    const float keep = hi ? v[i + n] : v[i], send = hi ? v[i] : v[i + n];
                    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
                }
            }
#pragma unroll
            for (int o = 16 >> H; o >= LPR; o >>= 1) {
#pragma unroll
                for (int i = 0; i < NV; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
            }
            // lanes with bit 16 clear hold sum e, their partners sum e^2, of coordinate q*4 + comp
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const float sq = __shfl_xor_sync(0xffffffffu, v[i], 16);
                if ((lane & (16 | BF_MASK)) == 0) {
                    const int comp = (((lane & 8) ? 2 : 0) | ((H >= 3 && (lane & 4)) ? 1 : 0)) + i;
                    const int j = q * 4 + comp;
                    fmrow[Fc + C + j] = 0.5f * (v[i] * v[i] - sq);      // models/deepfm.py:88-92
                    sum_e[b * (int64_t) K + j] = v[i];
                }
            }
        }
        __syncwarp();        // every lane is done with the slot before the next fill overwrites it
    }
}

// K2: one warp per (sample, multi-valued field)
template <int K>
__global__ void __launch_bounds__(256)
pooled_gather_fwd_kernel(const ctr_embed_cfg cfg, const float* __restrict__ table, const void* __restrict__ ids,
                         const int32_t* __restrict__ multi_start, const float* __restrict__ slot_weight,
                         float* __restrict__ x0, float* __restrict__ pooled_first, float* __restrict__ inv_cnt_emb,
                         float* __restrict__ inv_cnt_first) {
    constexpr int LPR = K / 4;
    constexpr int RPW = 32 / LPR;
    const int lane = threadIdx.x & 31;
    const int64_t item = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int Fm = cfg.Fm;
    if (item >= (int64_t) cfg.B * Fm) return;
    const int64_t b = item / Fm;
    const int j = (int) (item - b * Fm);
    const int r = lane / LPR, q = lane % LPR;
    const int s0 = multi_start[j], s1 = multi_start[j + 1];
    const int stride = cfg.table_stride;
    const bool has_first = cfg.has_first != 0, zero0 = cfg.zero_row0 != 0;
    const int64_t idrow = b * (int64_t) cfg.ids_ld + cfg.Fc;

    float4 acc = f4_zero();
    float accw = 0.f;
    int cnt_e = 0, cnt_w = 0;
    for (int t0 = s0; t0 < s1; t0 += RPW) {
        int t = t0 + r;
        bool ok = t < s1;
        int64_t rid = ok ? ctr_load_id(ids, cfg.id_bytes, idrow + t) : 0;
        const bool bad = ok && cfg.table_rows > 0 && (uint64_t) rid >= (uint64_t) cfg.table_rows;
        if (bad) ctr_flag_bad_id(cfg.status, cfg.Fc + t);
        bool live = ok && !bad && !(zero0 && rid == 0);
        const float* row = table + rid * (int64_t) stride;
        float4 e = live ? __ldg(reinterpret_cast<const float4*>(row + q * 4)) : (bad ? f4_nan() : f4_zero());
        float w = (live && has_first && q == 0) ? __ldg(row + K) : 0.f;
        float sw = (ok && slot_weight) ? slot_weight[b * (int64_t) cfg.M + t] : 1.f;
        // row sum over the k coordinates (models/deepfm_multi.py:84 axis_2_sum)
        float rs = (e.x + e.y) + (e.z + e.w);
#pragma unroll
        for (int o = 1; o < LPR; o <<= 1) rs += __shfl_xor_sync(0xffffffffu, rs, o);
        bool nz_e = cfg.count_rule == 0 ? (rs != 0.f) : live;
        bool nz_w = cfg.count_rule == 0 ? (w != 0.f) : live;
        cnt_e += __popc(__ballot_sync(0xffffffffu, ok && q == 0 && nz_e));
        cnt_w += __popc(__ballot_sync(0xffffffffu, ok && q == 0 && nz_w));
        acc.x = fmaf(e.x, sw, acc.x); acc.y = fmaf(e.y, sw, acc.y);
        acc.z = fmaf(e.z, sw, acc.z); acc.w = fmaf(e.w, sw, acc.w);
        accw = fmaf(w, sw, accw);
    }
#pragma unroll
    for (int o = LPR; o < 32; o <<= 1) {
        acc = f4_shfl_xor_add(acc, o);
        accw += __shfl_xor_sync(0xffffffffu, accw, o);
    }
    const bool mean = cfg.combiner == 0;
    if (r == 0) {
        float4 p = acc;
        if (mean) {
            if (cnt_e > 0) {
                float c = (float) cnt_e;
                p = make_float4(acc.x / c, acc.y / c, acc.z / c, acc.w / c);
            } else {
                p = f4_zero();   // tf.div_no_nan
            }
        }
        if (cfg.round_tf32) p = make_float4(ctr_round_tf32(p.x), ctr_round_tf32(p.y), ctr_round_tf32(p.z), ctr_round_tf32(p.w));
        *reinterpret_cast<float4*>(x0 + b * (int64_t) cfg.x0_ld + (cfg.Fc + j) * K + q * 4) = p;
    }
    if (lane == 0) {
        float pw = accw;
        if (mean) pw = cnt_w > 0 ? accw / (float) cnt_w : 0.f;
        pooled_first[item] = pw;
        inv_cnt_emb[item] = mean ? (cnt_e > 0 ? 1.f / (float) cnt_e : 0.f) : 1.f;
        inv_cnt_first[item] = mean ? (cnt_w > 0 ? 1.f / (float) cnt_w : 0.f) : 1.f;
    }
}

__global__ void scalar_gather_sum_kernel(const float* __restrict__ table, const void* __restrict__ ids, int id_bytes,
                                         int B, int W, int64_t rows, uint32_t* __restrict__ status, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t b = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= B) return;
    float acc = 0.f;
    for (int j = lane; j < W; j += 32) {
        const int64_t id = ctr_load_id(ids, id_bytes, b * (int64_t) W + j);
        if (rows > 0 && (uint64_t) id >= (uint64_t) rows) {
            ctr_flag_bad_id(status, j);
            acc = __int_as_float(0x7fc00000);
        } else {
            acc += __ldg(table + id);
        }
    }
    acc = ctr_warp_sum(acc);
    if (lane == 0) out[b] = acc;
}

int check_cfg(ctr_handle* h, const ctr_embed_cfg* c) {
    CTR_REQUIRE(h, c != nullptr, "embed cfg is null");
    CTR_REQUIRE(h, c->k == 4 || c->k == 8 || c->k == 16 || c->k == 32, "embedding_size %d unsupported (4/8/16/32)", c->k);
    CTR_REQUIRE(h, c->id_bytes == 4 || c->id_bytes == 8, "id_bytes must be 4 or 8");
    CTR_REQUIRE(h, c->B >= 0 && c->Fc >= 0 && c->Fm >= 0 && c->C >= 0 && c->V >= 0 && c->M >= 0, "negative sizes");
    CTR_REQUIRE(h, c->ids_ld >= c->Fc + c->M, "ids_ld < Fc + M");
    CTR_REQUIRE(h, c->table_stride >= c->k + (c->has_first ? 1 : 0) && c->table_stride % 4 == 0,
                "table_stride %d too small / not a multiple of 4", c->table_stride);
    CTR_REQUIRE(h, c->x0_ld % 4 == 0 && c->x0_ld >= (c->Fc + c->Fm) * c->k + c->C + c->V, "x0_ld %d invalid", c->x0_ld);
    CTR_REQUIRE(h, !c->use_fm || c->fm_ld >= c->Fc + c->Fm + c->C + c->k, "fm_ld too small");
    return CTR_OK;
}

}  // namespace

int ctr_check_embed_cfg(ctr_handle* h, const ctr_embed_cfg* c) { return check_cfg(h, c); }

extern "C" int ctr_fm_gather_fwd(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg, const float* table,
                                 const void* ids, int64_t id_offset, const float* cont, const float* vec,
                                 const float* cont_rows, const float* pooled_first, float* x0, float* fm,
                                 float* sum_e) {
    int rc = check_cfg(h, cfg);
    if (rc) return rc;
    CTR_REQUIRE(h, x0 && (cfg->Fc == 0 || (table && ids)), "fm_gather_fwd: null table/ids/x0");
    CTR_REQUIRE(h, cfg->C == 0 || cont, "fm_gather_fwd: cont is null");
    CTR_REQUIRE(h, cfg->V == 0 || vec, "fm_gather_fwd: vec is null");
    CTR_REQUIRE(h, !cfg->use_fm || (fm && sum_e), "fm_gather_fwd: fm/sum_e null with use_fm");
    CTR_REQUIRE(h, !(cfg->use_fm && cfg->fm_cont && cfg->C > 0) || cont_rows, "fm_gather_fwd: cont_rows null");
    if (cfg->B == 0) return CTR_OK;
    // fast path: the headline shapes (DeepFM / DNN / WDL with single-valued fields only)
    {
        const int lpr = cfg->k / 4, rpw = 32 / lpr;
        const bool fast = cfg->Fm == 0 && cfg->V == 0 && cfg->C <= 2 * rpw && cfg->C <= 32 &&
                          cfg->x0_ld - cfg->Fc * cfg->k <= 32 && cfg->k <= 32 && getenv("CTR_K1_GENERAL") == nullptr;
        // staged path (cp.async ring + TMA bulk store of the X0 row) unless CTR_K1=regs asks for the register-staged kernel
        // measured (profiles/r02b_k1_variants.txt, B = 64K, Zipf / uniform ids): register-staged 47.3 / 57.9 us, staged
        // 57.0 / 67.9 us whatever the ring depth (4, 6) or CTAs per SM (2, 3, 4) -- so neither kernel is bound by bytes in
        // flight; the staged kernel stays selectable (CTR_K1=staged) and covered by the parity tests
        static const char* k1_mode = getenv("CTR_K1");
        const bool staged = fast && cfg->Fc <= 32 && cfg->Fc > 0 &&
                            (reinterpret_cast<uintptr_t>(x0) & 15) == 0 && (reinterpret_cast<uintptr_t>(table) & 15) == 0 &&
                            (k1_mode && strcmp(k1_mode, "staged") == 0);
        if (staged) {
            static int sms = 0, ns_sel = 0, cta_cap = 0;
            if (!sms) {
                int dev = 0;
                cudaGetDevice(&dev);
                cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
                const char* e = getenv("CTR_K1_NS");
                ns_sel = (e && atoi(e) == 6) ? 6 : 4;
                e = getenv("CTR_K1_CTAS");
                cta_cap = e ? atoi(e) : 0;
            }
            static int occ_s[4][2][2][2];
            const int ki = cfg->k == 4 ? 0 : cfg->k == 8 ? 1 : cfg->k == 16 ? 2 : 3, fi = cfg->use_fm ? 1 : 0, ii = cfg->id_bytes == 8,
                      ni = ns_sel == 6;
            const size_t smem = (size_t) 8 * ns_sel * (cfg->Fc * cfg->k + 36) * 4;
            CTR_REQUIRE(h, smem <= 227 * 1024, "fm_gather_fwd: embedding block too wide for the staged kernel");
#define LAUNCH_STAGE(KK, FM_, I64, NS_)                                                                                \
    {                                                                                                                  \
        auto kern = fm_gather_stage_kernel<KK, FM_, I64, NS_>;                                                         \
        int& o = occ_s[ki][fi][ii][ni];                                                                                \
        if (!o) {                                                                                                      \
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);                       \
            int q = 0;                                                                                                 \
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&q, kern, 256, smem);                                        \
            o = q > 0 ? q : 1;                                                                                         \
        }                                                                                                              \
        const int per_sm = (cta_cap > 0 && cta_cap < o) ? cta_cap : o;                                                 \
        const int64_t want = ctr_cdiv((int64_t) cfg->B, 8), cap = (int64_t) sms * per_sm;                              \
        kern<<<(unsigned) (want < cap ? want : cap), 256, smem, ctr_st(s)>>>(*cfg, table, ids, id_offset, cont,        \
                                                                            cont_rows, x0, fm, sum_e);                 \
    }
#define LAUNCH_STAGE_N(KK, FM_, I64) \
    if (ni) LAUNCH_STAGE(KK, FM_, I64, 6) else LAUNCH_STAGE(KK, FM_, I64, 4)
#define LAUNCH_STAGE_K(KK)                                                           \
    if (cfg->use_fm) {                                                               \
        if (ii) { LAUNCH_STAGE_N(KK, true, true) } else { LAUNCH_STAGE_N(KK, true, false) }   \
    } else {                                                                         \
        if (ii) { LAUNCH_STAGE_N(KK, false, true) } else { LAUNCH_STAGE_N(KK, false, false) } \
    }
            switch (cfg->k) {
                case 4: LAUNCH_STAGE_K(4) break;
                case 8: LAUNCH_STAGE_K(8) break;
                case 16: LAUNCH_STAGE_K(16) break;
                default: LAUNCH_STAGE_K(32) break;
            }
#undef LAUNCH_STAGE_K
#undef LAUNCH_STAGE_N
#undef LAUNCH_STAGE
            CTR_CHECK_LAUNCH(h, "fm_gather_fwd(staged)");
            return CTR_OK;
        }
        if (fast) {
            static int occ[4][2][2];         // resident CTAs per SM of each instantiation, queried once
            static int sms = 0;
            if (!sms) {
                int dev = 0;
                cudaGetDevice(&dev);
                cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            }
            const int ki = cfg->k == 4 ? 0 : cfg->k == 8 ? 1 : cfg->k == 16 ? 2 : 3, fi = cfg->use_fm ? 1 : 0, ii = cfg->id_bytes == 8;
#define LAUNCH_FAST(KK, FM_, I64)                                                                                      \
    {                                                                                                                  \
        auto kern = fm_gather_fast_kernel<KK, FM_, I64>;                                                               \
        if (!occ[ki][fi][ii]) {                                                                                        \
            int o = 0;                                                                                                 \
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kern, 256, 0);                                           \
            occ[ki][fi][ii] = o > 0 ? o : 1;                                                                           \
        }                                                                                                              \
        const int64_t want = ctr_cdiv((int64_t) cfg->B, 8), cap = (int64_t) sms * occ[ki][fi][ii];                     \
        kern<<<(unsigned) (want < cap ? want : cap), 256, 0, ctr_st(s)>>>(*cfg, table, ids, id_offset, cont, cont_rows, \
                                                                         x0, fm, sum_e);                               \
    }
#define LAUNCH_FAST_K(KK)                                                    \
    if (cfg->use_fm) {                                                       \
        if (ii) LAUNCH_FAST(KK, true, true) else LAUNCH_FAST(KK, true, false)    \
    } else {                                                                 \
        if (ii) LAUNCH_FAST(KK, false, true) else LAUNCH_FAST(KK, false, false)  \
    }
            switch (cfg->k) {
                case 4: LAUNCH_FAST_K(4) break;
                case 8: LAUNCH_FAST_K(8) break;
                case 16: LAUNCH_FAST_K(16) break;
                default: LAUNCH_FAST_K(32) break;
            }
#undef LAUNCH_FAST_K
#undef LAUNCH_FAST
            CTR_CHECK_LAUNCH(h, "fm_gather_fwd(fast)");
            return CTR_OK;
        }
    }
    const unsigned blocks = (unsigned) ctr_cdiv((int64_t) cfg->B * 32, 256);
#define LAUNCH_K1(KK)                                                                                              \
    if (cfg->use_fm)                                                                                               \
        fm_gather_fwd_kernel<KK, true><<<blocks, 256, 0, ctr_st(s)>>>(*cfg, table, ids, id_offset, cont, vec,      \
                                                                      cont_rows, pooled_first, x0, fm, sum_e);     \
    else                                                                                                           \
        fm_gather_fwd_kernel<KK, false><<<blocks, 256, 0, ctr_st(s)>>>(*cfg, table, ids, id_offset, cont, vec,     \
                                                                       cont_rows, pooled_first, x0, fm, sum_e)
    switch (cfg->k) {
        case 4: LAUNCH_K1(4); break;
        case 8: LAUNCH_K1(8); break;
        case 16: LAUNCH_K1(16); break;
        default: LAUNCH_K1(32); break;
    }
#undef LAUNCH_K1
    CTR_CHECK_LAUNCH(h, "fm_gather_fwd");
    return CTR_OK;
}

extern "C" int ctr_pooled_gather_fwd(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg, const float* table,
                                     const void* ids, const int32_t* multi_start, const float* slot_weight, float* x0,
                                     float* pooled_first, float* inv_cnt_emb, float* inv_cnt_first) {
    int rc = check_cfg(h, cfg);
    if (rc) return rc;
    if (cfg->Fm == 0 || cfg->B == 0) return CTR_OK;
    CTR_REQUIRE(h, table && ids && multi_start && x0 && pooled_first && inv_cnt_emb && inv_cnt_first,
                "pooled_gather_fwd: null pointer");
    const int64_t items = (int64_t) cfg->B * cfg->Fm;
    const unsigned blocks = (unsigned) ctr_cdiv(items * 32, 256);
#define LAUNCH_K2(KK)                                                                                            \
    pooled_gather_fwd_kernel<KK><<<blocks, 256, 0, ctr_st(s)>>>(*cfg, table, ids, multi_start, slot_weight, x0,  \
                                                               pooled_first, inv_cnt_emb, inv_cnt_first)
    switch (cfg->k) {
        case 4: LAUNCH_K2(4); break;
        case 8: LAUNCH_K2(8); break;
        case 16: LAUNCH_K2(16); break;
        default: LAUNCH_K2(32); break;
    }
#undef LAUNCH_K2
    CTR_CHECK_LAUNCH(h, "pooled_gather_fwd");
    return CTR_OK;
}

extern "C" int ctr_scalar_gather_sum(ctr_handle* h, ctr_stream s, const float* table, const void* ids, int id_bytes,
                                     int B, int W, int64_t rows, uint32_t* status, float* out) {
    CTR_REQUIRE(h, table && ids && out && (id_bytes == 4 || id_bytes == 8) && B >= 0 && W >= 0,
                "scalar_gather_sum: bad args");
    if (B == 0) return CTR_OK;
    scalar_gather_sum_kernel<<<(unsigned) ctr_cdiv((int64_t) B * 32, 256), 256, 0, ctr_st(s)>>>(table, ids, id_bytes, B,
                                                                                               W, rows, status, out);
    CTR_CHECK_LAUNCH(h, "scalar_gather_sum");
    return CTR_OK;
}
