// K4: backward of the embedding side.  No per-lookup gradient tensor is ever materialised: the
// sorted lookup list is walked in fixed chunks and the gradient of each lookup is formed on the
// fly from dX0 (the MLP's input gradient), the saved FM sum and the head weights, accumulated in
// registers per run of equal row ids and written once per (row, chunk).  The Adam update then
// consumes those partials in chunk order, so the whole scatter-add is deterministic.
#include <stdlib.h>

#include "ctr_common.cuh"
#include "peer_route.cuh"

namespace {

constexpr int CH = CTR_BWD_CHUNK;
constexpr int LONG_RUN = 8;   // runs with more partials than this take the warp-cooperative path

// Layout of one gradient row (grad_stride floats): [0,K) embedding coordinates, [K] first-order weight, [K+1] D.
// For a single-valued lookup of row r by sample b the FM part of the gradient is dl_b * w2 * (s_b - e_r)
// (models/deepfm.py:88-92 differentiated).  e_r is the same for every lookup of the run, so the walker only adds
// dl_b * w2 * s_b and D += dl_b; the update kernel, which reads e_r anyway, subtracts w2 * e_r * D.  That removes the
// random re-read of the X0 tile (one of three 64-byte random reads per lookup).  Multi-valued slots enter the FM
// through their pooled vector, not the row, so they keep the direct form and add nothing to D.
template <int K, int UN = 4, int MINB = 3>
__global__ void __launch_bounds__(256, MINB)
embed_bwd_partials_kernel(const ctr_embed_cfg cfg, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ perm,
                          const uint32_t* __restrict__ uidx, const int64_t n, const int32_t* __restrict__ field_of_slot,
                          const float* __restrict__ x0, const float* __restrict__ dx0, const int64_t dx0_ld,
                          const float* __restrict__ sum_e, const float* __restrict__ dlogit,
                          const float* __restrict__ w_fm, const float* __restrict__ inv_cnt_emb,
                          const float* __restrict__ inv_cnt_first, const float* __restrict__ cont_grad,
                          const float* __restrict__ slot_weight, float* __restrict__ part) {
    constexpr int G = K / 4;   // lanes per chunk walker: lane q owns coordinates 4q..4q+3
    // UN = lookups in flight per walker (template parameter)
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t c = t / G;   // chunk index
    const int q = (int) (t % G);
    const int64_t base = c * CH;
    if (base >= n) return;
    const int cnt = (int) ((n - base) < CH ? (n - base) : CH);
    const int Fc = cfg.Fc, Fm = cfg.Fm, gs = cfg.grad_stride;
    const uint32_t ids_ld = (uint32_t) cfg.ids_ld;
    const bool use_fm = cfg.use_fm != 0, has_first = cfg.has_first != 0;
    const float* __restrict__ w1 = w_fm;                         // first-order head weights, per field
    float4 w2 = make_float4(0.f, 0.f, 0.f, 0.f);                 // second-order head weights of this lane's coordinates
    if (use_fm) {
        const float* w2p = w_fm + Fc + Fm + cfg.C + q * 4;
        w2 = make_float4(w2p[0], w2p[1], w2p[2], w2p[3]);
    }
    const uint32_t n_lookup = (uint32_t) cfg.B * ids_ld;

    uint32_t cur = CTR_PAD_KEY;
    int64_t cur_slot = 0;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    float accw = 0.f, accd = 0.f;

    auto flush = [&]() {
        float* o = part + cur_slot * gs;
        *reinterpret_cast<float4*>(o + q * 4) = acc;
        if (q == 0 && (has_first || use_fm)) *reinterpret_cast<float2*>(o + K) = make_float2(accw, accd);
    };

    for (int i0 = 0; i0 < cnt; i0 += UN) {
        uint32_t key[UN], u[UN];
        float4 gv[UN];
        float gw[UN], gd[UN];
#pragma unroll
        for (int k = 0; k < UN; ++k) {
            const int i = i0 + k;
            key[k] = CTR_PAD_KEY; u[k] = 0; gv[k] = make_float4(0.f, 0.f, 0.f, 0.f); gw[k] = 0.f; gd[k] = 0.f;
            if (i < cnt) {
                key[k] = __ldg(keys + base + i);
                if (key[k] != CTR_PAD_KEY) {
                    u[k] = __ldg(uidx + base + i);
                    const uint32_t p = __ldg(perm + base + i);
                    if (p >= n_lookup) {   // appended continuous-field row: gradient was reduced densely
                        const int cc = (int) (p - n_lookup);
                        gv[k] = *reinterpret_cast<const float4*>(cont_grad + cc * gs + q * 4);
                        gw[k] = has_first ? cont_grad[cc * gs + K] : 0.f;
                        continue;
                    }
                    const uint32_t b = p / ids_ld;
                    const int slot = (int) (p - b * ids_ld);
                    if (slot < Fc) {
                        float4 g = __ldg(reinterpret_cast<const float4*>(dx0 + (int64_t) b * dx0_ld + slot * K + q * 4));
                        if (use_fm) {
                            const float dl = __ldg(dlogit + b);
                            const float4 sv = __ldg(reinterpret_cast<const float4*>(sum_e + (int64_t) b * K + q * 4));
                            g.x = fmaf(dl * w2.x, sv.x, g.x); g.y = fmaf(dl * w2.y, sv.y, g.y);
                            g.z = fmaf(dl * w2.z, sv.z, g.z); g.w = fmaf(dl * w2.w, sv.w, g.w);
                            gw[k] = dl * __ldg(w1 + slot);
                            gd[k] = dl;
                        }
                        gv[k] = g;
                    } else {
                        const int jm = field_of_slot[slot - Fc];
                        const int fld = Fc + jm;
                        float sc_e = inv_cnt_emb[(int64_t) b * Fm + jm], sc_w = inv_cnt_first[(int64_t) b * Fm + jm];
                        if (slot_weight) {   // weighted pooling (models/dnn_multi_textline.py:96-101): d row = w[b,t] * upstream / count
                            const float sw = __ldg(slot_weight + (int64_t) b * cfg.M + (slot - Fc));
                            sc_e *= sw; sc_w *= sw;
                        }
                        float4 g = __ldg(reinterpret_cast<const float4*>(dx0 + (int64_t) b * dx0_ld + fld * K + q * 4));
                        if (use_fm) {
                            const float dl = __ldg(dlogit + b);
                            const float4 sv = __ldg(reinterpret_cast<const float4*>(sum_e + (int64_t) b * K + q * 4));
                            const float4 xv = __ldg(reinterpret_cast<const float4*>(x0 + (int64_t) b * cfg.x0_ld + fld * K + q * 4));
                            g.x = fmaf(dl * w2.x, sv.x - xv.x, g.x); g.y = fmaf(dl * w2.y, sv.y - xv.y, g.y);
                            g.z = fmaf(dl * w2.z, sv.z - xv.z, g.z); g.w = fmaf(dl * w2.w, sv.w - xv.w, g.w);
                            gw[k] = sc_w * dl * __ldg(w1 + fld);
                        }
                        gv[k] = make_float4(sc_e * g.x, sc_e * g.y, sc_e * g.z, sc_e * g.w);
                    }
                }
            }
        }
#pragma unroll
        for (int k = 0; k < UN; ++k) {
            if (i0 + k < cnt) {
                if (key[k] != cur) {
                    if (cur != CTR_PAD_KEY) flush();
                    cur = key[k];
                    cur_slot = (int64_t) u[k] + c;
                    acc = make_float4(0.f, 0.f, 0.f, 0.f); accw = 0.f; accd = 0.f;
                }
                acc.x += gv[k].x; acc.y += gv[k].y; acc.z += gv[k].z; acc.w += gv[k].w;
                accw += gw[k];
                accd += gd[k];
            }
        }
    }
    if (cur != CTR_PAD_KEY) flush();
}

// sum of the partials of run u, coordinate j, in chunk order
__device__ __forceinline__ float sum_run_partials(const float* __restrict__ part, const uint32_t* __restrict__ seg_start,
                                                  int64_t u, int j, int gs) {
    const uint32_t s = seg_start[u], e = seg_start[u + 1];
    const int64_t c0 = s / CH, c1 = (e - 1) / CH;
    const float* p = part + (u + c0) * gs + j;
    const int64_t np = c1 - c0 + 1;
    if (np == 1) return p[0];
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    int64_t i = 0;
    for (; i + 4 <= np; i += 4) {
        a0 += p[(i + 0) * gs];
        a1 += p[(i + 1) * gs];
        a2 += p[(i + 2) * gs];
        a3 += p[(i + 3) * gs];
    }
    for (; i < np; ++i) a0 += p[i * gs];
    return (a0 + a1) + (a2 + a3);
}

// grads[u, :width] = sum of the partial rows of run u, in chunk order.  One thread per (run, 4 columns) with 128-bit
// accesses; a run spread over more than LONG_RUN chunks (a hot row: thousands of lookups) is summed by the whole warp
// -- lane l takes partial rows l, l + 32, ..., then a fixed shuffle tree -- so one thread never walks hundreds of rows.
// PUSH: the summed row goes straight into the owner's receive buffer over NVLink (the gradient all-to-all, fused);
// runs past the routed list (the padding key) are dropped.
template <bool PUSH>
__global__ void __launch_bounds__(256)
embed_bwd_finalize_kernel(const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ n_unique, int64_t n_max,
                          int width, int gs, const float* __restrict__ part, float* __restrict__ grads, const PeerRoute rt) {
    int64_t U = *n_unique;
    if (U > n_max) U = n_max;
    if (PUSH) {
        const int64_t routed = route_total(rt, (int64_t) rt.seg_end[rt.world - 1]);     // runs past it = the padding key
        if (U > routed) U = routed;
    }
    auto out_row = [&](int64_t u) -> float* { return PUSH ? route_row(rt, (uint32_t) u, gs) : grads + u * gs; };
    const int nq = (width + 3) >> 2;                 // column quads per row
    const int64_t total = U * nq;
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t) gridDim.x * blockDim.x;
    for (int64_t t0 = (int64_t) blockIdx.x * blockDim.x + (threadIdx.x & ~31); t0 < total; t0 += stride) {   // warp-uniform
        const int64_t t = t0 + lane;
        const bool valid = t < total;
        const int64_t u = valid ? t / nq : 0;
        const int q = valid ? (int) (t - u * nq) : 0;
        int64_t c0 = 0, np = 0;
        if (valid) {
            const uint32_t s0 = __ldg(seg_start + u), s1 = __ldg(seg_start + u + 1);
            c0 = s0 / CH;
            np = (int64_t) ((s1 - 1) / CH) - c0 + 1;
        }
        const bool lng = valid && np > LONG_RUN;
        if (valid && !lng) {
            const float* p = part + (u + c0) * gs + q * 4;
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int64_t i = 0; i < np; ++i) {
                const float4 x = __ldg(reinterpret_cast<const float4*>(p + i * gs));
                a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
            }
            *reinterpret_cast<float4*>(out_row(u) + q * 4) = a;
        }
        unsigned todo = __ballot_sync(0xffffffffu, lng && q == 0);     // the run's first quad leads
        while (todo) {
            const int l = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t uu = __shfl_sync(0xffffffffu, u, l), cc = __shfl_sync(0xffffffffu, c0, l), nn = __shfl_sync(0xffffffffu, np, l);
            const float4* base = reinterpret_cast<const float4*>(part + (uu + cc) * gs);
            const int gs4 = gs >> 2;
            for (int qq = 0; qq < nq; ++qq) {
                float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int64_t i = lane; i < nn; i += 32) {
                    const float4 x = __ldg(base + i * gs4 + qq);
                    a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
                }
                a.x = ctr_warp_sum(a.x); a.y = ctr_warp_sum(a.y); a.z = ctr_warp_sum(a.z); a.w = ctr_warp_sum(a.w);
                if (lane == 0) *reinterpret_cast<float4*>(out_row(uu) + qq * 4) = a;
            }
        }
    }
}

// One element of the optimizer step, shared by every update kernel of this file.
//   kind 0  TF1 Adam (adam.py, epsilon-hat form):  m <- b1 m + c1 g ; v <- b2 v + c2 g^2 ; th -= lr_t m / (sqrt(v) + eps)
//   kind 1  TF1 FTRL (ftrl.py -> ApplyFtrl / SparseApplyFtrl with the defaults lr_power = -0.5, l1 = l2 = 0;
//           test_model/wdl.py:136), slot m = `accum`, slot v = `linear`, hp.lr_t = the (decayed) learning rate:
//             accum' = accum + g^2 ; linear += g - (sqrt(accum') - sqrt(accum)) / lr * th ; th = -linear / (sqrt(accum') / lr)
//           (th = 0 where linear == 0: a looked-up weight whose gradient has always been zero is SET to zero, as the
//           TF kernel does)
__device__ __forceinline__ void opt_update(float& th, float& m, float& v, const float g, const ctr_adam_hp& hp) {
    if (hp.kind == 1) {
        // sqrt(accum') - sqrt(accum) is written as g^2 / (sqrt(accum') + sqrt(accum)): the same number without the
        // cancellation (accum = 0.1, g^2 ~ 1e-8: in float32 the difference of the two roots keeps 1-2 digits, and it is
        // multiplied by th / lr.  The reference's own file run in float32 lands 1.5e-4 from its float64 run for this reason,
        // tests/golden/fp32floor_wdl_ftrl.npz)
        const float gg = g * g;
        const float acc_new = m + gg;
        const float sq_new = sqrtf(acc_new);
        const float lin = v + g - gg / (sq_new + sqrtf(m)) / hp.lr_t * th;
        m = acc_new;
        v = lin;
        th = lin != 0.f ? -lin / (sq_new / hp.lr_t) : 0.f;
        return;
    }
    m = hp.beta1 * m + hp.c1 * g;
    v = hp.beta2 * v + hp.c2 * g * g;
    th -= hp.lr_t * m / (sqrtf(v) + hp.eps);
}

// generic width (any table): one thread per (row, coordinate)
__global__ void __launch_bounds__(256)
adam_rows_kernel(const uint32_t* __restrict__ row_ids, const uint32_t* __restrict__ n_rows, int64_t n_max, uint32_t mask,
                 const float* __restrict__ grads, const uint32_t* __restrict__ seg_start, int gs, int width,
                 float* __restrict__ table, float* __restrict__ m, float* __restrict__ v, int stride, int64_t table_rows,
                 uint8_t* __restrict__ touched, uint32_t* __restrict__ long_list, const float* __restrict__ fm_w2, int emb_k,
                 const ctr_adam_hp hp) {
    if (ctr_guard_tripped(hp.guard)) return;
    int64_t U = n_rows ? (int64_t) *n_rows : n_max;
    if (U > n_max) U = n_max;
    const int64_t total = U * width;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t u = t / width;
        const int j = (int) (t - u * width);
        const uint32_t key = row_ids[u];
        if (key == CTR_PAD_KEY) continue;
        const int64_t row = (int64_t) (key & mask);
        if (row >= table_rows) continue;   // defensive: never write outside the table
        if (seg_start && long_list) {
            // runs spread over many chunks (hot rows) go to the warp-cooperative kernel
            const int64_t np = (int64_t) ((seg_start[u + 1] - 1) / CH) - (int64_t) (seg_start[u] / CH) + 1;
            if (np > LONG_RUN) {
                if (j == 0) long_list[1 + atomicAdd(long_list, 1u)] = (uint32_t) u;
                continue;
            }
        }
        float g = seg_start ? sum_run_partials(grads, seg_start, u, j, gs) : grads[u * gs + j];
        const int64_t o = row * stride + j;
        if (fm_w2 && j < emb_k) {          // the e-term of the FM gradient (see embed_bwd_partials_kernel)
            const float D = seg_start ? sum_run_partials(grads, seg_start, u, emb_k + 1, gs) : grads[u * gs + emb_k + 1];
            g = fmaf(-fm_w2[j] * table[o], D, g);
        }
        float th_ = table[o], mm = m[o], vv = v[o];
        opt_update(th_, mm, vv, g, hp);
        m[o] = mm;
        v[o] = vv;
        table[o] = th_;
        if (j == 0 && touched) touched[row] = 1;
    }
}

__device__ __forceinline__ void adam4(float4& th, float4& mm, float4& vv, const float4 g, const ctr_adam_hp& hp) {
    opt_update(th.x, mm.x, vv.x, g.x, hp); opt_update(th.y, mm.y, vv.y, g.y, hp);
    opt_update(th.z, mm.z, vv.z, g.z, hp); opt_update(th.w, mm.w, vv.w, g.w, hp);
}

// embedding tables (row = K coordinates [+ first-order weight]): K/4 lanes per row, 128-bit accesses, every load of a
// row independent of the others.  The arithmetic per element is the generic kernel's, in the same order.
template <int K, int MINB = 4>
__global__ void __launch_bounds__(256, MINB)
adam_rows_vec_kernel(const uint32_t* __restrict__ row_ids, const uint32_t* __restrict__ n_rows, int64_t n_max, uint32_t mask,
                     const float* __restrict__ grads, const uint32_t* __restrict__ seg_start, int gs, int has_first,
                     float* __restrict__ table, float* __restrict__ m, float* __restrict__ v, int stride, int64_t table_rows,
                     uint8_t* __restrict__ touched, uint32_t* __restrict__ long_list, const float* __restrict__ fm_w2,
                     const uint32_t* __restrict__ perm, const ctr_adam_hp hp) {
    // perm != NULL (owner side of the sharded step): run u is the contributions grads[perm[i]], i in [seg_start[u],
    // seg_start[u+1]) -- one row per requesting peer, in peer order -- summed here instead of by ctr_segment_sum_rows
    constexpr int G = K / 4;
    if (ctr_guard_tripped(hp.guard)) return;
    int64_t U = n_rows ? (int64_t) *n_rows : n_max;
    if (U > n_max) U = n_max;
    const int64_t total = U * G;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t u = t / G;
        const int q = (int) (t - u * G);
        const uint32_t key = __ldg(row_ids + u);
        if (key == CTR_PAD_KEY) continue;
        const int64_t row = (int64_t) (key & mask);
        if (row >= table_rows) continue;
        const float* p = grads + u * gs;
        int np = 1;
        uint32_t r0 = 0;
        if (perm) {
            r0 = __ldg(seg_start + u);
            np = (int) (__ldg(seg_start + u + 1) - r0);
        } else if (seg_start) {
            const uint32_t s0 = __ldg(seg_start + u), s1 = __ldg(seg_start + u + 1);
            const int64_t c0 = s0 / CH;
            const int64_t npl = (int64_t) ((s1 - 1) / CH) - c0 + 1;
            if (long_list && npl > LONG_RUN) {
                if (q == 0) long_list[1 + atomicAdd(long_list, 1u)] = (uint32_t) u;
                continue;
            }
            np = (int) npl;
            p = grads + (u + c0) * gs;
        }
        const int64_t o = row * stride + q * 4;
        float4 th = *reinterpret_cast<const float4*>(table + o);
        float4 mm = *reinterpret_cast<const float4*>(m + o);
        float4 vv = *reinterpret_cast<const float4*>(v + o);
        float th1 = 0.f, m1 = 0.f, v1 = 0.f;
        const bool first = has_first && q == 0;
        if (first) { th1 = table[row * stride + K]; m1 = m[row * stride + K]; v1 = v[row * stride + K]; }
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        float g1 = 0.f, D = 0.f;
        for (int i = 0; i < np; ++i) {       // partials in chunk order / peer contributions in run order
            const float* pi = perm ? grads + (int64_t) __ldg(perm + r0 + i) * gs : p + (int64_t) i * gs;
            const float4 a = __ldg(reinterpret_cast<const float4*>(pi + q * 4));
            g.x += a.x; g.y += a.y; g.z += a.z; g.w += a.w;
            if (first) g1 += __ldg(pi + K);
            if (fm_w2) D += __ldg(pi + K + 1);
        }
        if (fm_w2) {
            const float4 w2 = make_float4(__ldg(fm_w2 + q * 4), __ldg(fm_w2 + q * 4 + 1), __ldg(fm_w2 + q * 4 + 2), __ldg(fm_w2 + q * 4 + 3));
            g.x = fmaf(-w2.x * th.x, D, g.x); g.y = fmaf(-w2.y * th.y, D, g.y);
            g.z = fmaf(-w2.z * th.z, D, g.z); g.w = fmaf(-w2.w * th.w, D, g.w);
        }
        adam4(th, mm, vv, g, hp);
        *reinterpret_cast<float4*>(m + o) = mm;
        *reinterpret_cast<float4*>(v + o) = vv;
        *reinterpret_cast<float4*>(table + o) = th;
        if (first) {
            opt_update(th1, m1, v1, g1, hp);
            m[row * stride + K] = m1;
            v[row * stride + K] = v1;
            table[row * stride + K] = th1;
        }
        if (q == 0 && touched) touched[row] = 1;
    }
}

// hot rows: one warp per run.  Lane l sums partial rows l, l+32, ... for ALL coordinates at once
// (128-bit loads of the contiguous [np, gs] block), then a fixed shuffle tree: deterministic.
constexpr int MAX_GS4 = 9;   // grad_stride / 4 for k = 32 with a first-order column
__global__ void __launch_bounds__(256)
adam_rows_long_kernel(const uint32_t* __restrict__ row_ids, uint32_t mask, const float* __restrict__ part,
                      const uint32_t* __restrict__ seg_start, int gs, int width, float* __restrict__ table,
                      float* __restrict__ m, float* __restrict__ v, int stride, int64_t table_rows,
                      uint8_t* __restrict__ touched, const uint32_t* __restrict__ long_list,
                      const float* __restrict__ fm_w2, int emb_k, const ctr_adam_hp hp) {
    const int lane = threadIdx.x & 31;
    const int gs4 = gs >> 2;
    if (ctr_guard_tripped(hp.guard)) return;
    const uint32_t nlong = long_list[0];
    const int64_t nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    for (int64_t w = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < nlong; w += nwarps) {
        const int64_t u = long_list[1 + w];
        const uint32_t s = seg_start[u], e = seg_start[u + 1];
        const int64_t c0 = s / CH, np = (int64_t) ((e - 1) / CH) - c0 + 1;
        const float4* base = reinterpret_cast<const float4*>(part + (u + c0) * gs);
        float4 acc[MAX_GS4];
#pragma unroll
        for (int q = 0; q < MAX_GS4; ++q) acc[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int64_t i = lane; i < np; i += 32) {
#pragma unroll
            for (int q = 0; q < MAX_GS4; ++q) {
                if (q < gs4) {
                    const float4 t = base[i * gs4 + q];
                    acc[q].x += t.x; acc[q].y += t.y; acc[q].z += t.z; acc[q].w += t.w;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < MAX_GS4; ++q) {
            if (q < gs4) {
                acc[q].x = ctr_warp_sum(acc[q].x); acc[q].y = ctr_warp_sum(acc[q].y);
                acc[q].z = ctr_warp_sum(acc[q].z); acc[q].w = ctr_warp_sum(acc[q].w);
            }
        }
        const uint32_t key = row_ids[u];
        const int64_t row = (int64_t) (key & mask);
        if (key == CTR_PAD_KEY || row >= table_rows) continue;
        auto pick = [&](int j) {
            float g = 0.f;
#pragma unroll
            for (int q = 0; q < MAX_GS4; ++q) {
                if (q == (j >> 2)) {
                    const int r = j & 3;
                    g = r == 0 ? acc[q].x : (r == 1 ? acc[q].y : (r == 2 ? acc[q].z : acc[q].w));
                }
            }
            return g;
        };
        const float D = fm_w2 ? pick(emb_k + 1) : 0.f;
        // lane j applies Adam to coordinate j (j + 32 for very wide rows)
        for (int j = lane; j < width; j += 32) {
            float g = pick(j);
            const int64_t o = row * stride + j;
            if (fm_w2 && j < emb_k) g = fmaf(-fm_w2[j] * table[o], D, g);
            float th_ = table[o], mm = m[o], vv = v[o];
            opt_update(th_, mm, vv, g, hp);
            m[o] = mm;
            v[o] = vv;
            table[o] = th_;
        }
        if (lane == 0 && touched) touched[row] = 1;
    }
}

__global__ void __launch_bounds__(256)
adam_dense_kernel(float* __restrict__ theta, float* __restrict__ m, float* __restrict__ v, const float* __restrict__ g,
                  int64_t n, float l2, float l1, uint8_t* __restrict__ skip, int skip_group, const ctr_adam_hp hp) {
    if (ctr_guard_tripped(hp.guard)) return;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        if (skip && skip[i / skip_group]) continue;
        const float th = theta[i];
        float gi = g ? g[i] : 0.f;
        if (l2 != 0.f) gi = fmaf(l2, th, gi);
        if (l1 != 0.f) gi += l1 * (th > 0.f ? 1.f : (th < 0.f ? -1.f : 0.f));
        float nt_ = th, mm = m[i], vv = v[i];
        opt_update(nt_, mm, vv, gi, hp);
        m[i] = mm;
        v[i] = vv;
        theta[i] = nt_;
    }
}

// whole dense side of one step in ONE launch: TF1 Adam over the flat parameter buffer (three regulariser groups by
// position) and, for the tower weights, the K-major and TF32-rounded copies the tcgen05 GEMMs of the next step read.
// Replaces 2-3 adam_dense + 2 launches per layer (8 launches of a few microseconds each at config 4).
__global__ void __launch_bounds__(256)
dense_update_kernel(float* __restrict__ theta, float* __restrict__ m, float* __restrict__ v, const float* __restrict__ g,
                    const ctr_dense_layout lay, const ctr_adam_hp hp) {
    if (ctr_guard_tripped(hp.guard)) return;
    const int64_t n = lay.group_end[2];
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const float th = theta[i];
        float gi = g[i];
        if (i >= lay.group_end[1]) gi += lay.reg_scale * (th > 0.f ? 1.f : (th < 0.f ? -1.f : 0.f));     // l1 group
        else if (i >= lay.group_end[0]) gi = fmaf(lay.reg_scale, th, gi);                                 // l2 group
        float nt = th, mm = m[i], vv = v[i];
        opt_update(nt, mm, vv, gi, hp);
        m[i] = mm;
        v[i] = vv;
        theta[i] = nt;
#pragma unroll 1
        for (int l = 0; l < lay.n_layers; ++l) {
            const int64_t r = i - lay.w_off[l];
            if (r >= 0 && r < (int64_t) lay.K[l] * lay.N[l]) {
                const float q = ctr_round_tf32(nt);
                if (lay.w_rounded[l]) lay.w_rounded[l][r] = q;
                if (lay.w_kmajor[l]) {
                    const int row = (int) (r / lay.N[l]), col = (int) (r - (int64_t) row * lay.N[l]);
                    lay.w_kmajor[l][(int64_t) col * lay.ld_kmajor[l] + row] = q;
                }
                break;
            }
        }
    }
}

__global__ void clear_flags_kernel(uint8_t* __restrict__ f, int64_t n) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) f[i] = 0;
}

// gradient of the continuous-field rows, pass 1.  K lanes own the K coordinates of one sample slot, a warp walks
// 32/K samples at a time (coalesced sum_e rows, broadcast cont / dlogit), every lane keeps one accumulator per
// field; fields are taken CONT_FT at a time.  Fixed sample -> (block, warp, slot) assignment and fixed reduction
// trees: deterministic.  Block partials go to ws[block][c*(K+1) + j] (j == K: the first-order weight).
constexpr int CONT_FT = 16;
template <int K>
__global__ void __launch_bounds__(256)
cont_bwd_pass1(const ctr_embed_cfg cfg, const float* __restrict__ cont, const float* __restrict__ cont_rows,
               const float* __restrict__ sum_e, const float* __restrict__ dlogit, const float* __restrict__ w_fm,
               float* __restrict__ ws) {
    constexpr int SPW = 32 / K;                 // sample slots per warp
    constexpr int W = K + 1;
    __shared__ float red[8][CONT_FT * (K + 1)];
    const int C = cfg.C, npairs = C * W;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int j = lane % K, slot = lane / K;
    const int64_t per = ((int64_t) cfg.B + gridDim.x - 1) / gridDim.x;
    const int64_t lo = per * blockIdx.x, hi = (lo + per < cfg.B) ? lo + per : cfg.B;
    const float* w1c = w_fm + cfg.Fc + cfg.Fm;
    const float w2j = w_fm[cfg.Fc + cfg.Fm + C + j];
    const bool has_first = cfg.has_first != 0;
    for (int c0 = 0; c0 < C; c0 += CONT_FT) {
        float th[CONT_FT], acc[CONT_FT], accw[CONT_FT];
#pragma unroll
        for (int t = 0; t < CONT_FT; ++t) {
            th[t] = (c0 + t < C) ? cont_rows[(int64_t) (c0 + t) * cfg.table_stride + j] : 0.f;
            acc[t] = 0.f;
            accw[t] = 0.f;
        }
        for (int64_t b = lo + (int64_t) w * SPW + slot; b < hi; b += 8 * SPW) {
            const float dl = dlogit[b];
            const float se = sum_e[b * K + j];
            const float* xr = cont + b * C + c0;
#pragma unroll
            for (int t = 0; t < CONT_FT; ++t) {
                if (c0 + t < C) {
                    const float x = xr[t];
                    const float xd = x * dl;
                    acc[t] = fmaf(xd * w2j, se - x * th[t], acc[t]);
                    accw[t] += xd;
                }
            }
        }
        // sample slots of the warp, then the 8 warps of the block
#pragma unroll
        for (int t = 0; t < CONT_FT; ++t) {
#pragma unroll
            for (int o = K; o < 32; o <<= 1) {
                acc[t] += __shfl_xor_sync(0xffffffffu, acc[t], o);
                accw[t] += __shfl_xor_sync(0xffffffffu, accw[t], o);
            }
        }
        if (slot == 0) {
#pragma unroll
            for (int t = 0; t < CONT_FT; ++t) {
                red[w][t * W + j] = acc[t];
                if (j == 0) red[w][t * W + K] = accw[t];
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < CONT_FT * W; i += blockDim.x) {
            const int t = i / W, jj = i - t * W;
            if (c0 + t < C) {
                float a = 0.f;
#pragma unroll
                for (int ww = 0; ww < 8; ++ww) a += red[ww][i];
                if (jj == K) a = has_first ? a * w1c[c0 + t] : 0.f;
                ws[(int64_t) blockIdx.x * npairs + (c0 + t) * W + jj] = a;
            }
        }
        __syncthreads();
    }
}

// one warp per (field, coordinate): lane-strided partial sums, then the fixed shuffle tree
__global__ void cont_bwd_pass2(const float* __restrict__ ws, int nblk, int C, int K, int gs, float* __restrict__ out) {
    const int W = K + 1, npairs = C * W;
    const int pr = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (pr >= npairs) return;
    float acc = 0.f;
    for (int b = lane; b < nblk; b += 32) acc += ws[(int64_t) b * npairs + pr];
    acc = ctr_warp_sum(acc);
    const int c = pr / W, j = pr - c * W;
    if (lane == 0) out[c * gs + j] = acc;
}

// wide part of Wide&Deep (models/wdl.py:240-246): d wdl_weights[id] += sum over the lookups (b, j) of that id of
// dlogit[b].  The lookups arrive sorted by id (ctr_sort_ids + ctr_unique_segments).  Three shapes of run, each with a fixed
// order of additions (the result never changes from call to call):
//   <= 64 lookups     8 lanes walk the run with a fixed stride, then a fixed shuffle tree
//   <= SSA_HUGE       the whole warp, eight independent loads per lane in flight
//   longer (the hottest ids of a Zipf-distributed field: tens of thousands of lookups, and ONE warp walking them was
//                     0.24 ms of config 3): cut into pieces of SSA_PIECE lookups that a second launch sums one block per
//                     piece, and a third adds up in piece order.  Pieces and runs are listed in the caller's workspace;
//                     a run that does not fit the lists is summed by its warp after all.
constexpr uint32_t SSA_HUGE = 2048, SSA_PIECE = 4096;
constexpr int SSA_MAX_RUNS = 1024, SSA_MAX_PIECES = 8192;
struct SsaWs {
    uint32_t n_pieces, n_runs, pad0, pad1;
    uint32_t run_key[SSA_MAX_RUNS], run_first[SSA_MAX_RUNS], run_np[SSA_MAX_RUNS];
    uint32_t piece_lo[SSA_MAX_PIECES], piece_hi[SSA_MAX_PIECES];
    float piece_sum[SSA_MAX_PIECES];
};

__global__ void __launch_bounds__(256)
scalar_segment_add_kernel(const uint32_t* __restrict__ uniq, const uint32_t* __restrict__ seg_start,
                          const uint32_t* __restrict__ n_unique, int64_t n_max, const uint32_t* __restrict__ perm,
                          const float* __restrict__ src, uint32_t ld, int64_t dst_rows, float* __restrict__ dst,
                          SsaWs* __restrict__ ws) {
    int64_t U = *n_unique;
    if (U > n_max) U = n_max;
    const int lane = threadIdx.x & 31, sub = lane & 7;
    const int64_t groups = ((int64_t) gridDim.x * blockDim.x) >> 3;
    // every lane of a warp runs the same number of iterations (the shuffles below are warp-wide)
    const int64_t g0 = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const int64_t iters = (U + groups - 1) / groups;
    for (int64_t it = 0; it < iters; ++it) {
        const int64_t u = g0 + it * groups;
        float acc = 0.f;
        uint32_t key = CTR_PAD_KEY, s = 0, e = 0;
        if (u < U) {
            key = uniq[u];
            s = seg_start[u];
            e = seg_start[u + 1];
        }
        const bool live = key != CTR_PAD_KEY;
        const bool is_long = live && (e - s) > 64u;
        int listed = 0;
        if (ws && is_long && sub == 0 && (e - s) > SSA_HUGE && (int64_t) key < dst_rows) {
            const uint32_t np = (e - s + SSA_PIECE - 1) / SSA_PIECE;
            const uint32_t r = atomicAdd(&ws->n_runs, 1u);
            if (r < (uint32_t) SSA_MAX_RUNS) {
                const uint32_t first = atomicAdd(&ws->n_pieces, np);
                if (first + np <= (uint32_t) SSA_MAX_PIECES) {
                    ws->run_key[r] = key; ws->run_first[r] = first; ws->run_np[r] = np;
                    for (uint32_t p = 0; p < np; ++p) {
                        ws->piece_lo[first + p] = s + p * SSA_PIECE;
                        ws->piece_hi[first + p] = (s + (p + 1) * SSA_PIECE < e) ? s + (p + 1) * SSA_PIECE : e;
                    }
                    listed = 1;
                } else {
                    // no room for its pieces: an empty run entry, and the slots it reserved below the cap become empty pieces
                    // (every slot under min(n_pieces, cap) is then either a real piece or a defined empty one); the warp
                    // sums the run below
                    ws->run_np[r] = 0;
                    for (uint32_t p = first; p < first + np && p < (uint32_t) SSA_MAX_PIECES; ++p) ws->piece_lo[p] = ws->piece_hi[p] = 0u;
                }
            }
        }
        listed = __shfl_sync(0xffffffffu, listed, lane & ~7);
        if (live && !is_long)
            for (uint32_t i = s + sub; i < e; i += 8) acc += src[perm[i] / ld];
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        acc += __shfl_xor_sync(0xffffffffu, acc, 4);
        unsigned todo = __ballot_sync(0xffffffffu, is_long && !listed && sub == 0);
        while (todo) {
            const int owner = __ffs(todo) - 1;
            todo &= todo - 1;
            const uint32_t ls = __shfl_sync(0xffffffffu, s, owner), le = __shfl_sync(0xffffffffu, e, owner);
            float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            uint32_t i = ls + lane;
            for (; i + 7 * 32 < le; i += 8 * 32) {
#pragma unroll
                for (int j = 0; j < 8; ++j) a[j] += src[perm[i + j * 32] / ld];
            }
            for (; i < le; i += 32) a[0] += src[perm[i] / ld];
            float t = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
            if (lane == owner) acc = t;
        }
        if (sub == 0 && live && !listed && (int64_t) key < dst_rows) dst[key] += acc;
    }
}

// one block per listed piece: threads stride over its lookups, four loads in flight each, fixed reduction tree
__global__ void __launch_bounds__(256)
scalar_segment_pieces_kernel(SsaWs* __restrict__ ws, const uint32_t* __restrict__ perm, const float* __restrict__ src, uint32_t ld) {
    __shared__ float part[8];
    uint32_t np = ws->n_pieces;
    if (np > (uint32_t) SSA_MAX_PIECES) np = SSA_MAX_PIECES;
    for (uint32_t p = blockIdx.x; p < np; p += gridDim.x) {
        const uint32_t lo = ws->piece_lo[p], hi = ws->piece_hi[p];
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        uint32_t i = lo + threadIdx.x;
        for (; i + 3 * 256 < hi; i += 4 * 256) {
            a0 += src[perm[i] / ld]; a1 += src[perm[i + 256] / ld];
            a2 += src[perm[i + 512] / ld]; a3 += src[perm[i + 768] / ld];
        }
        for (; i < hi; i += 256) a0 += src[perm[i] / ld];
        float t = (a0 + a1) + (a2 + a3);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = t;
        __syncthreads();
        if (threadIdx.x == 0) {
            float r = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) r += part[w];
            ws->piece_sum[p] = r;
        }
        __syncthreads();
    }
}

// one thread per listed run: its pieces in order; then the counters go back to zero for the next call
__global__ void __launch_bounds__(256)
scalar_segment_runs_kernel(SsaWs* __restrict__ ws, int64_t dst_rows, float* __restrict__ dst) {
    uint32_t nr = ws->n_runs;
    if (nr > (uint32_t) SSA_MAX_RUNS) nr = SSA_MAX_RUNS;
    for (uint32_t r = threadIdx.x; r < nr; r += blockDim.x) {
        const uint32_t np = ws->run_np[r], first = ws->run_first[r], key = ws->run_key[r];
        if (np == 0) continue;                    // listed without room for its pieces: its warp summed it
        float t = 0.f;
        for (uint32_t p = 0; p < np; ++p) t += ws->piece_sum[first + p];
        if ((int64_t) key < dst_rows) dst[key] += t;
    }
    __syncthreads();
    if (threadIdx.x == 0) { ws->n_pieces = 0; ws->n_runs = 0; }
}

// l2 term over GATHERED scalar weights (models/wdl_textline.py:264-269,284-286: the regulariser sees the [B, W+H] lookup
// result, not the table).  A row looked up c times carries c * w^2 / 2 in the loss and scale * c * w in its gradient.
__global__ void __launch_bounds__(256)
scalar_gathered_l2_runs_kernel(const uint32_t* __restrict__ uniq, const uint32_t* __restrict__ seg_start,
                               const uint32_t* __restrict__ n_unique, int64_t n_max, const float* __restrict__ w, float scale,
                               int64_t rows, float* __restrict__ grad, float* __restrict__ sq) {
    int64_t U = *n_unique;
    if (U > n_max) U = n_max;
    for (int64_t u = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; u < n_max; u += (int64_t) gridDim.x * blockDim.x) {
        float q = 0.f;
        if (u < U) {
            const uint32_t key = uniq[u];
            if (key != CTR_PAD_KEY && (int64_t) key < rows) {
                const float c = (float) (seg_start[u + 1] - seg_start[u]), x = w[key];
                grad[key] += scale * c * x;       // keys are unique: one writer per row
                q = c * x * x;
            }
        }
        sq[u] = q;
    }
}

__global__ void __launch_bounds__(256)
scalar_gathered_l2_dense_kernel(const float* __restrict__ w, float scale, int64_t lo, int n, float count,
                                float* __restrict__ grad, float* __restrict__ sq) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const float x = w[lo + j];
    grad[lo + j] += scale * count * x;
    sq[j] = count * x * x;
}

__global__ void __launch_bounds__(256)
segment_sum_rows_kernel(const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ n_unique, int64_t n_max,
                        const uint32_t* __restrict__ perm, const float* __restrict__ src, int src_stride, int width,
                        float* __restrict__ dst, int dst_stride) {
    int64_t U = *n_unique;
    if (U > n_max) U = n_max;
    const int64_t total = U * width;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t u = t / width;
        const int j = (int) (t - u * width);
        float acc = 0.f;
        for (uint32_t i = seg_start[u], e = seg_start[u + 1]; i < e; ++i) acc += src[(int64_t) perm[i] * src_stride + j];
        dst[u * dst_stride + j] = acc;
    }
}

// 128-bit form: one thread per (run, 4 columns); strides are multiples of 4 floats, so the pad columns of a row ride along
__global__ void __launch_bounds__(256)
segment_sum_rows_vec_kernel(const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ n_unique, int64_t n_max,
                            const uint32_t* __restrict__ perm, const float* __restrict__ src, int src_stride, int nq,
                            float* __restrict__ dst, int dst_stride) {
    int64_t U = *n_unique;
    if (U > n_max) U = n_max;
    const int64_t total = U * nq;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t u = t / nq;
        const int q = (int) (t - u * nq);
        const uint32_t s0 = __ldg(seg_start + u), s1 = __ldg(seg_start + u + 1);
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t i = s0;
        for (; i + 2 <= s1; i += 2) {        // two contributions in flight; the additions stay in run order
            const float4 x = __ldg(reinterpret_cast<const float4*>(src + (int64_t) __ldg(perm + i) * src_stride) + q);
            const float4 y = __ldg(reinterpret_cast<const float4*>(src + (int64_t) __ldg(perm + i + 1) * src_stride) + q);
            a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
            a.x += y.x; a.y += y.y; a.z += y.z; a.w += y.w;
        }
        if (i < s1) {
            const float4 x = __ldg(reinterpret_cast<const float4*>(src + (int64_t) __ldg(perm + i) * src_stride) + q);
            a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
        }
        *(reinterpret_cast<float4*>(dst + u * dst_stride) + q) = a;
    }
}

inline int cont_blocks(ctr_handle* h, int B) {
    int64_t nb = ctr_cdiv(B > 0 ? B : 1, 128);
    int64_t cap = (int64_t) h->sm_count * 2;
    return (int) (nb < cap ? nb : cap);
}

}  // namespace

int ctr_check_embed_cfg(ctr_handle* h, const ctr_embed_cfg* c);   // embed_fwd.cu

extern "C" int ctr_embed_bwd_partials_weighted(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg,
                                               const uint32_t* sorted_keys, const uint32_t* perm, const uint32_t* uidx,
                                               int64_t n, const int32_t* field_of_slot, const float* x0, const float* dx0,
                                               int64_t dx0_ld, const float* sum_e, const float* dlogit, const float* w_fm,
                                               const float* inv_cnt_emb, const float* inv_cnt_first, const float* cont_grad,
                                               const float* slot_weight, float* part);

extern "C" int ctr_embed_bwd_partials(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg, const uint32_t* sorted_keys,
                                      const uint32_t* perm, const uint32_t* uidx, int64_t n, const int32_t* field_of_slot,
                                      const float* x0, const float* dx0, int64_t dx0_ld, const float* sum_e,
                                      const float* dlogit, const float* w_fm, const float* inv_cnt_emb,
                                      const float* inv_cnt_first, const float* cont_grad, float* part) {
    return ctr_embed_bwd_partials_weighted(h, s, cfg, sorted_keys, perm, uidx, n, field_of_slot, x0, dx0, dx0_ld, sum_e, dlogit,
                                           w_fm, inv_cnt_emb, inv_cnt_first, cont_grad, nullptr, part);
}

extern "C" int ctr_embed_bwd_partials_weighted(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg,
                                               const uint32_t* sorted_keys, const uint32_t* perm, const uint32_t* uidx,
                                               int64_t n, const int32_t* field_of_slot, const float* x0, const float* dx0,
                                               int64_t dx0_ld, const float* sum_e, const float* dlogit, const float* w_fm,
                                               const float* inv_cnt_emb, const float* inv_cnt_first, const float* cont_grad,
                                               const float* slot_weight, float* part) {
    int rc = ctr_check_embed_cfg(h, cfg);
    if (rc) return rc;
    CTR_REQUIRE(h, sorted_keys && perm && uidx && dx0 && part && n >= 0, "embed_bwd_partials: null pointer");
    CTR_REQUIRE(h, cfg->grad_stride >= cfg->k + ((cfg->has_first || cfg->use_fm) ? 2 : 0) && cfg->grad_stride % 4 == 0,
                "embed_bwd_partials: grad_stride must be a multiple of 4, >= k + 2 with first-order / FM columns");
    CTR_REQUIRE(h, dx0_ld % 4 == 0 && cfg->x0_ld % 4 == 0 && ((reinterpret_cast<uintptr_t>(dx0) | reinterpret_cast<uintptr_t>(part) |
                       reinterpret_cast<uintptr_t>(sum_e) | reinterpret_cast<uintptr_t>(x0) | reinterpret_cast<uintptr_t>(cont_grad)) & 15) == 0,
                "embed_bwd_partials: dx0 / x0 / sum_e / part / cont_grad must be 16-byte aligned with leading dimensions % 4 == 0");
    CTR_REQUIRE(h, !cfg->use_fm || (x0 && sum_e && dlogit && w_fm), "embed_bwd_partials: FM inputs null");
    CTR_REQUIRE(h, cfg->M == 0 || (field_of_slot && inv_cnt_emb && inv_cnt_first), "embed_bwd_partials: multi inputs null");
    CTR_REQUIRE(h, dx0_ld >= (int64_t) (cfg->Fc + cfg->Fm) * cfg->k, "embed_bwd_partials: dx0_ld too small");
    CTR_REQUIRE(h, !slot_weight || cfg->M > 0, "embed_bwd_partials: slot_weight without multi-valued slots");
    if (n == 0) return CTR_OK;
    const int64_t chunks = ctr_cdiv(n, CH);
    const int64_t threads = chunks * (cfg->k / 4);
    const unsigned blocks = (unsigned) ctr_cdiv(threads, 256);
#define LAUNCH_BP(KK, ...)                                                                                             \
    embed_bwd_partials_kernel<KK, ##__VA_ARGS__><<<blocks, 256, 0, ctr_st(s)>>>(*cfg, sorted_keys, perm, uidx, n, field_of_slot, x0, dx0, \
                                                                dx0_ld, sum_e, dlogit, w_fm, inv_cnt_emb, inv_cnt_first,   \
                                                                cont_grad, slot_weight, part)
    // k = 16 (config 4): ONE lookup in flight per walker and 6 CTAs per SM (40 registers).  The 833 CTAs of a 64K batch then
    // run as a single wave; with 4 lookups in flight (72 registers, 3 CTAs per SM) it was 1.9 waves of long serial walks:
    // 85.7 -> 71.3 us in the step (profiles/r02l_partials_variants.txt).  CTR_BP_VARIANT=8 selects the old shape.
    static int variant = -1;
    if (variant < 0) {
        const char* e = getenv("CTR_BP_VARIANT");
        variant = e ? atoi(e) : 9;
    }
    switch (cfg->k) {
        case 4: LAUNCH_BP(4); break;
        case 8: LAUNCH_BP(8); break;
        case 16:
            if (variant == 8) LAUNCH_BP(16, 4, 3);
            else if (variant == 4) LAUNCH_BP(16, 2, 4);
            else LAUNCH_BP(16, 1, 6);
            break;
        default: LAUNCH_BP(32); break;
    }
#undef LAUNCH_BP
    CTR_CHECK_LAUNCH(h, "embed_bwd_partials");
    return CTR_OK;
}

static inline unsigned grid_for(ctr_handle* h, int64_t total) {
    int64_t blocks = ctr_cdiv(total > 0 ? total : 1, 256);
    int64_t cap = (int64_t) h->sm_count * 16;
    return (unsigned) (blocks < cap ? blocks : cap);
}

extern "C" int ctr_embed_bwd_finalize(ctr_handle* h, ctr_stream s, const uint32_t* seg_start, const uint32_t* n_unique,
                                      int64_t n_max, int width, int grad_stride, const float* part, float* grads) {
    CTR_REQUIRE(h, seg_start && n_unique && part && grads && width > 0 && grad_stride >= width && n_max >= 0,
                "embed_bwd_finalize: bad args");
    if (n_max == 0) return CTR_OK;
    CTR_REQUIRE(h, grad_stride % 4 == 0 && ((reinterpret_cast<uintptr_t>(part) | reinterpret_cast<uintptr_t>(grads)) & 15) == 0,
                "embed_bwd_finalize: grad_stride must be a multiple of 4 and part / grads 16-byte aligned");
    embed_bwd_finalize_kernel<false><<<grid_for(h, n_max * ((width + 3) / 4)), 256, 0, ctr_st(s)>>>(
        seg_start, n_unique, n_max, width, grad_stride, part, grads, PeerRoute{});
    CTR_CHECK_LAUNCH(h, "embed_bwd_finalize");
    return CTR_OK;
}

extern "C" int ctr_embed_bwd_finalize_push(ctr_handle* h, ctr_stream s, const uint32_t* seg_start, const uint32_t* n_unique,
                                           int64_t n_max, int width, int grad_stride, const float* part,
                                           const ctr_peer_route* rt) {
    CTR_REQUIRE(h, seg_start && n_unique && part && rt && width > 0 && grad_stride >= width && n_max >= 0,
                "embed_bwd_finalize_push: bad args");
    CTR_REQUIRE(h, rt->world >= 1 && rt->world <= CTR_MAX_PEERS && grad_stride % 4 == 0 && rt->dst_off % 16 == 0 &&
                       (reinterpret_cast<uintptr_t>(part) & 15) == 0,
                "embed_bwd_finalize_push: grad_stride must be a multiple of 4, buffers 16-byte aligned");
    if (!rt->dyn) {           // host-side route: the list length is known here
        const int64_t U = rt->seg_end[rt->world - 1];
        if (U == 0) return CTR_OK;
        if (U < n_max) n_max = U;
    }
    if (n_max == 0) return CTR_OK;
    embed_bwd_finalize_kernel<true><<<grid_for(h, n_max * ((width + 3) / 4)), 256, 0, ctr_st(s)>>>(
        seg_start, n_unique, n_max, width, grad_stride, part, nullptr, make_route(rt));
    CTR_CHECK_LAUNCH(h, "embed_bwd_finalize_push");
    return CTR_OK;
}

// a run with more than LONG_RUN partials holds more than (LONG_RUN - 1) * CH lookups
extern "C" int64_t ctr_adam_rows_ws_bytes(ctr_handle*, int64_t n_max) {
    return 4 * (n_max / ((LONG_RUN - 1) * CH) + 4) + 256;
}

extern "C" int ctr_adam_rows(ctr_handle* h, ctr_stream s, const uint32_t* row_ids, const uint32_t* n_rows, int64_t n_max,
                             int local_bits, const float* grads, const uint32_t* seg_start, int grad_stride, int width,
                             float* table, float* m, float* v, int table_stride, int64_t table_rows,
                             uint8_t* touched, void* ws, int64_t ws_bytes, const float* fm_w2, int emb_k,
                             const ctr_adam_hp* hp) {
    CTR_REQUIRE(h, row_ids && grads && table && m && v && hp, "adam_rows: null pointer");
    CTR_REQUIRE(h, (hp->kind == 1 ? hp->lr_t > 0.f : (hp->c1 > 0.f && hp->c2 > 0.f)) && (hp->kind == 0 || hp->kind == 1), "adam_rows: ctr_adam_hp.c1 / c2 (= 1 - beta, from the exact betas) must be set");
    CTR_REQUIRE(h, width > 0 && grad_stride >= width && table_stride >= width && n_max >= 0 && table_rows > 0,
                "adam_rows: bad sizes");
    CTR_REQUIRE(h, !fm_w2 || (emb_k > 0 && emb_k <= width && grad_stride >= emb_k + 2), "adam_rows: fm_w2 needs grad_stride >= emb_k + 2");
    if (n_max == 0) return CTR_OK;
    uint32_t mask = (local_bits <= 0 || local_bits >= 32) ? 0xFFFFFFFFu : ((1u << local_bits) - 1u);
    uint32_t* long_list = nullptr;
    if (seg_start) {
        CTR_REQUIRE(h, grad_stride % 4 == 0 && grad_stride / 4 <= MAX_GS4, "adam_rows: grad_stride must be a multiple of 4, <= 36");
        if (!ws || ws_bytes < ctr_adam_rows_ws_bytes(h, n_max)) return ctr_fail(h, CTR_E_WORKSPACE, "adam_rows: ws too small");
        long_list = (uint32_t*) ws;
        CTR_CUDA(h, cudaMemsetAsync(long_list, 0, 4, ctr_st(s)));
    }
    // embedding-table shape: vectorised kernel
    const bool aligned = ((reinterpret_cast<uintptr_t>(table) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
                           reinterpret_cast<uintptr_t>(grads)) & 15) == 0;
    const bool vec = emb_k > 0 && (emb_k == 4 || emb_k == 8 || emb_k == 16 || emb_k == 32) && (width == emb_k || width == emb_k + 1) &&
                     table_stride % 4 == 0 && grad_stride % 4 == 0 && aligned && (width == emb_k || grad_stride >= emb_k + 1);
    if (vec) {
        const int has_first = width == emb_k + 1;
        // 64 registers, 4 CTAs per SM.  More CTAs per SM (48 / 40 / 32 registers: spills) and persistent grids were measured
        // and are slower: 93.6 / 98.4 / 114.8 us against 83.7 (profiles/r02l_adam_rows_variants.txt)
        const unsigned grid = grid_for(h, n_max * (emb_k / 4));
#define LAUNCH_AV(KK)                                                                                                         \
    adam_rows_vec_kernel<KK><<<grid, 256, 0, ctr_st(s)>>>(row_ids, n_rows, n_max, mask, grads, seg_start, grad_stride, has_first, \
                                                          table, m, v, table_stride, table_rows, touched, long_list, fm_w2, nullptr, *hp)
        switch (emb_k) {
            case 4: LAUNCH_AV(4); break;
            case 8: LAUNCH_AV(8); break;
            case 16: LAUNCH_AV(16); break;
            default: LAUNCH_AV(32); break;
        }
#undef LAUNCH_AV
    } else {
        adam_rows_kernel<<<grid_for(h, n_max * width), 256, 0, ctr_st(s)>>>(row_ids, n_rows, n_max, mask, grads, seg_start,
                                                                             grad_stride, width, table, m, v, table_stride,
                                                                             table_rows, touched, long_list, fm_w2, emb_k, *hp);
    }
    CTR_CHECK_LAUNCH(h, "adam_rows");
    if (long_list) {
        adam_rows_long_kernel<<<(unsigned) h->sm_count * 4, 256, 0, ctr_st(s)>>>(row_ids, mask, grads, seg_start, grad_stride,
                                                                                 width, table, m, v, table_stride, table_rows,
                                                                                 touched, long_list, fm_w2, emb_k, *hp);
        CTR_CHECK_LAUNCH(h, "adam_rows_long");
    }
    return CTR_OK;
}

extern "C" int ctr_adam_rows_gathered(ctr_handle* h, ctr_stream s, const uint32_t* row_ids, const uint32_t* n_rows, int64_t n_max,
                                      int local_bits, const float* src, const uint32_t* perm, const uint32_t* seg_start,
                                      int src_stride, int width, float* table, float* m, float* v, int table_stride,
                                      int64_t table_rows, uint8_t* touched, const float* fm_w2, int emb_k, const ctr_adam_hp* hp) {
    CTR_REQUIRE(h, row_ids && n_rows && src && perm && seg_start && table && m && v && hp, "adam_rows_gathered: null pointer");
    CTR_REQUIRE(h, (hp->kind == 1 ? hp->lr_t > 0.f : (hp->c1 > 0.f && hp->c2 > 0.f)) && (hp->kind == 0 || hp->kind == 1), "adam_rows_gathered: ctr_adam_hp.c1 / c2 (= 1 - beta, from the exact betas) must be set");
    CTR_REQUIRE(h, n_max >= 0 && table_rows > 0 && (emb_k == 4 || emb_k == 8 || emb_k == 16 || emb_k == 32) &&
                       (width == emb_k || width == emb_k + 1) && table_stride % 4 == 0 && table_stride >= width &&
                       src_stride % 4 == 0 && src_stride >= width + (fm_w2 ? (emb_k + 2 - width) : 0),
                "adam_rows_gathered: embedding-row shapes only (k = 4/8/16/32, strides multiples of 4)");
    CTR_REQUIRE(h, ((reinterpret_cast<uintptr_t>(table) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
                     reinterpret_cast<uintptr_t>(src)) & 15) == 0, "adam_rows_gathered: buffers must be 16-byte aligned");
    if (n_max == 0) return CTR_OK;
    const uint32_t mask = (local_bits <= 0 || local_bits >= 32) ? 0xFFFFFFFFu : ((1u << local_bits) - 1u);
    const int has_first = width == emb_k + 1;
    const unsigned grid = (unsigned) ctr_cdiv(n_max * (emb_k / 4), 256);
#define LAUNCH_AG(KK)                                                                                                          \
    adam_rows_vec_kernel<KK><<<grid, 256, 0, ctr_st(s)>>>(row_ids, n_rows, n_max, mask, src, seg_start, src_stride, has_first,    \
                                                          table, m, v, table_stride, table_rows, touched, nullptr, fm_w2, perm, *hp)
    switch (emb_k) {
        case 4: LAUNCH_AG(4); break;
        case 8: LAUNCH_AG(8); break;
        case 16: LAUNCH_AG(16); break;
        default: LAUNCH_AG(32); break;
    }
#undef LAUNCH_AG
    CTR_CHECK_LAUNCH(h, "adam_rows_gathered");
    return CTR_OK;
}

extern "C" int ctr_adam_dense(ctr_handle* h, ctr_stream s, float* theta, float* m, float* v, const float* g, int64_t n,
                              float l2, float l1, uint8_t* skip, int skip_group, const ctr_adam_hp* hp) {
    CTR_REQUIRE(h, theta && m && v && hp && n >= 0 && (!skip || skip_group > 0), "adam_dense: bad args");
    CTR_REQUIRE(h, (hp->kind == 1 ? hp->lr_t > 0.f : (hp->c1 > 0.f && hp->c2 > 0.f)) && (hp->kind == 0 || hp->kind == 1), "adam_dense: ctr_adam_hp.c1 / c2 (= 1 - beta, from the exact betas) must be set");
    if (n == 0) return CTR_OK;
    adam_dense_kernel<<<grid_for(h, n), 256, 0, ctr_st(s)>>>(theta, m, v, g, n, l2, l1, skip, skip_group, *hp);
    CTR_CHECK_LAUNCH(h, "adam_dense");
    if (skip) {
        int64_t nf = ctr_cdiv(n, skip_group);
        clear_flags_kernel<<<grid_for(h, nf), 256, 0, ctr_st(s)>>>(skip, nf);
        CTR_CHECK_LAUNCH(h, "clear_flags");
    }
    return CTR_OK;
}

extern "C" int ctr_dense_update(ctr_handle* h, ctr_stream s, float* theta, float* m, float* v, const float* g,
                                const ctr_dense_layout* lay, const ctr_adam_hp* hp) {
    CTR_REQUIRE(h, theta && m && v && g && lay && hp, "dense_update: null pointer");
    CTR_REQUIRE(h, (hp->kind == 1 ? hp->lr_t > 0.f : (hp->c1 > 0.f && hp->c2 > 0.f)) && (hp->kind == 0 || hp->kind == 1), "dense_update: ctr_adam_hp.c1 / c2 (= 1 - beta, from the exact betas) must be set");
    CTR_REQUIRE(h, lay->group_end[0] >= 0 && lay->group_end[0] <= lay->group_end[1] && lay->group_end[1] <= lay->group_end[2] &&
                       lay->n_layers >= 0 && lay->n_layers <= CTR_MAX_LAYERS, "dense_update: bad layout");
    for (int l = 0; l < lay->n_layers; ++l)
        CTR_REQUIRE(h, lay->K[l] > 0 && lay->N[l] > 0 && lay->w_off[l] >= 0 &&
                           lay->w_off[l] + (int64_t) lay->K[l] * lay->N[l] <= lay->group_end[2] &&
                           (!lay->w_kmajor[l] || lay->ld_kmajor[l] >= lay->K[l]), "dense_update: bad layer %d", l);
    if (lay->group_end[2] == 0) return CTR_OK;
    dense_update_kernel<<<grid_for(h, lay->group_end[2]), 256, 0, ctr_st(s)>>>(theta, m, v, g, *lay, *hp);
    CTR_CHECK_LAUNCH(h, "dense_update");
    return CTR_OK;
}

extern "C" int64_t ctr_cont_bwd_ws_bytes(ctr_handle* h, const ctr_embed_cfg* cfg) {
    if (!h || !cfg) return 0;
    return (int64_t) cont_blocks(h, cfg->B) * cfg->C * (cfg->k + 1) * 4 + 256;
}

extern "C" int ctr_fm_cont_bwd(ctr_handle* h, ctr_stream s, const ctr_embed_cfg* cfg, const float* cont,
                               const float* cont_rows, const float* sum_e, const float* dlogit, const float* w_fm,
                               float* cont_grad, void* ws, int64_t ws_bytes) {
    int rc = ctr_check_embed_cfg(h, cfg);
    if (rc) return rc;
    if (cfg->C == 0 || cfg->B == 0) return CTR_OK;
    CTR_REQUIRE(h, cont && cont_rows && sum_e && dlogit && w_fm && cont_grad && ws, "fm_cont_bwd: null pointer");
    CTR_REQUIRE(h, cfg->grad_stride >= cfg->k + 1, "fm_cont_bwd: grad_stride too small");
    if (ws_bytes < ctr_cont_bwd_ws_bytes(h, cfg)) return ctr_fail(h, CTR_E_WORKSPACE, "fm_cont_bwd: ws too small");
    const int nblk = cont_blocks(h, cfg->B);
    switch (cfg->k) {
        case 4: cont_bwd_pass1<4><<<nblk, 256, 0, ctr_st(s)>>>(*cfg, cont, cont_rows, sum_e, dlogit, w_fm, (float*) ws); break;
        case 8: cont_bwd_pass1<8><<<nblk, 256, 0, ctr_st(s)>>>(*cfg, cont, cont_rows, sum_e, dlogit, w_fm, (float*) ws); break;
        case 16: cont_bwd_pass1<16><<<nblk, 256, 0, ctr_st(s)>>>(*cfg, cont, cont_rows, sum_e, dlogit, w_fm, (float*) ws); break;
        default: cont_bwd_pass1<32><<<nblk, 256, 0, ctr_st(s)>>>(*cfg, cont, cont_rows, sum_e, dlogit, w_fm, (float*) ws); break;
    }
    CTR_CHECK_LAUNCH(h, "cont_bwd_pass1");
    const int npairs = cfg->C * (cfg->k + 1);
    cont_bwd_pass2<<<(unsigned) ctr_cdiv((int64_t) npairs * 32, 256), 256, 0, ctr_st(s)>>>((const float*) ws, nblk, cfg->C, cfg->k,
                                                                          cfg->grad_stride, cont_grad);
    CTR_CHECK_LAUNCH(h, "cont_bwd_pass2");
    return CTR_OK;
}

extern "C" int64_t ctr_scalar_segment_add_ws_bytes(ctr_handle*) { return (int64_t) ((sizeof(SsaWs) + 255) / 256 * 256); }

extern "C" int ctr_scalar_segment_add(ctr_handle* h, ctr_stream s, const uint32_t* uniq, const uint32_t* seg_start,
                                      const uint32_t* n_unique, int64_t n_max, const uint32_t* perm, const float* src,
                                      int ld, int64_t dst_rows, float* dst, void* ws, int64_t ws_bytes) {
    CTR_REQUIRE(h, uniq && seg_start && n_unique && perm && src && dst && ld > 0 && dst_rows > 0 && n_max >= 0,
                "scalar_segment_add: bad args");
    if (n_max == 0) return CTR_OK;
    // ws (optional, ctr_scalar_segment_add_ws_bytes; its first 16 bytes zero on the first call, the call leaves them zero):
    // with it, runs of more than 2048 lookups are summed by many blocks instead of one warp
    SsaWs* w = (ws && ws_bytes >= (int64_t) sizeof(SsaWs)) ? (SsaWs*) ws : nullptr;
    scalar_segment_add_kernel<<<grid_for(h, n_max * 8), 256, 0, ctr_st(s)>>>(uniq, seg_start, n_unique, n_max, perm, src,
                                                                             (uint32_t) ld, dst_rows, dst, w);
    CTR_CHECK_LAUNCH(h, "scalar_segment_add");
    if (w) {
        scalar_segment_pieces_kernel<<<h->sm_count * 2, 256, 0, ctr_st(s)>>>(w, perm, src, (uint32_t) ld);
        scalar_segment_runs_kernel<<<1, 256, 0, ctr_st(s)>>>(w, dst_rows, dst);
        CTR_CHECK_LAUNCH(h, "scalar_segment_add (long runs)");
    }
    return CTR_OK;
}

extern "C" int ctr_scalar_gathered_l2(ctr_handle* h, ctr_stream s, const uint32_t* uniq, const uint32_t* seg_start,
                                      const uint32_t* n_unique, int64_t n_max, const float* w, float scale, int64_t rows,
                                      int64_t dense_lo, int dense_n, float dense_count, float* grad, float* sq) {
    CTR_REQUIRE(h, uniq && seg_start && n_unique && w && grad && sq && rows > 0 && n_max >= 0 && dense_n >= 0 &&
                       dense_lo >= 0 && dense_lo + dense_n <= rows, "scalar_gathered_l2: bad args");
    if (n_max > 0) {
        scalar_gathered_l2_runs_kernel<<<grid_for(h, n_max), 256, 0, ctr_st(s)>>>(uniq, seg_start, n_unique, n_max, w, scale,
                                                                                 rows, grad, sq);
        CTR_CHECK_LAUNCH(h, "scalar_gathered_l2_runs");
    }
    if (dense_n > 0) {
        // second launch on the same stream: a wide id may alias a dense row (quirk Q4), so the two never write concurrently
        scalar_gathered_l2_dense_kernel<<<(unsigned) ctr_cdiv((int64_t) dense_n, 256), 256, 0, ctr_st(s)>>>(
            w, scale, dense_lo, dense_n, dense_count, grad, sq + n_max);
        CTR_CHECK_LAUNCH(h, "scalar_gathered_l2_dense");
    }
    return CTR_OK;
}

extern "C" int ctr_segment_sum_rows(ctr_handle* h, ctr_stream s, const uint32_t* seg_start, const uint32_t* n_unique,
                                    int64_t n_max, const uint32_t* perm, const float* src, int src_stride, int width,
                                    float* dst, int dst_stride) {
    CTR_REQUIRE(h, seg_start && n_unique && perm && src && dst && width > 0 && src_stride >= width && dst_stride >= width &&
                       n_max >= 0, "segment_sum_rows: bad args");
    if (n_max == 0) return CTR_OK;
    const int nq = (width + 3) / 4;
    if (src_stride % 4 == 0 && dst_stride % 4 == 0 && nq * 4 <= src_stride && nq * 4 <= dst_stride &&
        ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
        int64_t blocks = ctr_cdiv(n_max * nq, 256);      // no cap: a thread's work is one short dependent chain
        segment_sum_rows_vec_kernel<<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(seg_start, n_unique, n_max, perm, src, src_stride,
                                                                              nq, dst, dst_stride);
        CTR_CHECK_LAUNCH(h, "segment_sum_rows_vec");
        return CTR_OK;
    }
    segment_sum_rows_kernel<<<grid_for(h, n_max * width), 256, 0, ctr_st(s)>>>(seg_start, n_unique, n_max, perm, src,
                                                                              src_stride, width, dst, dst_stride);
    CTR_CHECK_LAUNCH(h, "segment_sum_rows");
    return CTR_OK;
}
