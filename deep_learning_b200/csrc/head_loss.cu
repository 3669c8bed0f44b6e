// K5: output head (concat + [.,1] projection + sigmoid), tf.losses.log_loss and its gradient,
// plus the deterministic column reductions that give head-weight / bias gradients.
#include <stdlib.h>

#include "ctr_common.cuh"

namespace {

// per-sample loss and d loss / d logit (before the 1 / B_global factor) for score p = sigmoid(z), label y
//   flags & CTR_HEAD_SCE_ON_SCORE == 0: tf.losses.log_loss (models/deepfm.py:151-152):  -y log(p+e) - (1-y) log(1-p+e)
//   flags & CTR_HEAD_SCE_ON_SCORE:      test_model/wdl.py:119-121 feeds the ALREADY-sigmoided value to
//       tf.nn.sigmoid_cross_entropy_with_logits: with x = p in (0,1):  x - x y + log(1 + exp(-x)); d/dx = sigmoid(x) - y,
//       and dx/dz = p (1 - p)
__device__ __forceinline__ void head_loss(float p, float y, int flags, float eps, float& loss, float& dl_dz) {
    if (flags & CTR_HEAD_SCE_ON_SCORE) {
        loss = p - p * y + log1pf(expf(-p));
        dl_dz = (1.f / (1.f + expf(-p)) - y) * p * (1.f - p);
    } else {
        loss = -y * logf(p + eps) - (1.f - y) * logf(1.f - p + eps);
        dl_dz = (-y / (p + eps) + (1.f - y) / (1.f - p + eps)) * p * (1.f - p);
    }
}

constexpr float LOG_LOSS_EPS = 1e-7f;   // tf.losses.log_loss default epsilon

// One warp walks HEAD_SPW consecutive samples; lane l owns columns l, l+32, ... of the deep row (and of the fm row),
// so the row stays in registers between the dot product, the ReLU-gradient write and the head-weight gradient
//   d w_deep[j] = sum_b dlogit[b] * deep[b,j],  d w_fm[j] = sum_b dlogit[b] * fm[b,j],  d bias = sum_b dlogit[b]
// (models/deepfm.py:130,143 differentiated), which leave the kernel as one partial row per block: no separate
// pass over the activations.  All sums run in a fixed order (deterministic).
constexpr int HEAD_WARPS = 8;
constexpr int HEAD_SPW = 16;                       // samples per warp
constexpr int HEAD_SPB = HEAD_WARPS * HEAD_SPW;    // samples per block
constexpr int HEAD_MAXT = 16;                      // deep columns per lane  (H <= 512)
constexpr int HEAD_MAXF = 4;                       // fm columns per lane    (n_fm <= 128)

__global__ void __launch_bounds__(256)
head_fwd_kernel(int B, float inv_bg, const float* __restrict__ fm, int64_t fm_ld, int n_fm, const float* __restrict__ w_fm,
                const float* __restrict__ deep, int64_t deep_ld, int H, const float* __restrict__ w_deep,
                const float* __restrict__ extra, const float* __restrict__ bias, const float* __restrict__ label,
                float* __restrict__ logit, float* __restrict__ score, float* __restrict__ dlogit,
                float* __restrict__ dz_last, int64_t lddz, int round_tf32, float* __restrict__ loss_part,
                float* __restrict__ colpart, int flags) {
    const bool gate = (flags & CTR_HEAD_NO_RELU_GATE) == 0;      // dz_last passes through the tower's last ReLU
    extern __shared__ float sm[];                  // [HEAD_WARPS][H + n_fm + 1] when colpart, else [HEAD_WARPS]
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int ncol = H + n_fm + 1;
    float wd[HEAD_MAXT], wf[HEAD_MAXF], gd[HEAD_MAXT], gf[HEAD_MAXF];
#pragma unroll
    for (int t = 0; t < HEAD_MAXT; ++t) {
        const int j = lane + 32 * t;
        wd[t] = j < H ? w_deep[j] : 0.f;
        gd[t] = 0.f;
    }
#pragma unroll
    for (int t = 0; t < HEAD_MAXF; ++t) {
        const int j = lane + 32 * t;
        wf[t] = j < n_fm ? w_fm[j] : 0.f;
        gf[t] = 0.f;
    }
    float gb = 0.f, lsum = 0.f;
    const float bs = bias ? bias[0] : 0.f;
    const int64_t b0 = (int64_t) blockIdx.x * HEAD_SPB + (int64_t) w * HEAD_SPW;
    for (int i = 0; i < HEAD_SPW; ++i) {
        const int64_t b = b0 + i;
        if (b >= B) break;                         // uniform across the warp
        float rd[HEAD_MAXT], rf[HEAD_MAXF];
        float acc = 0.f;
        const float* r = deep + b * deep_ld;
#pragma unroll
        for (int t = 0; t < HEAD_MAXT; ++t) {
            const int j = lane + 32 * t;
            rd[t] = j < H ? r[j] : 0.f;
            acc = fmaf(rd[t], wd[t], acc);
        }
        const float* rfm = fm + b * fm_ld;
#pragma unroll
        for (int t = 0; t < HEAD_MAXF; ++t) {
            const int j = lane + 32 * t;
            rf[t] = j < n_fm ? rfm[j] : 0.f;
            acc = fmaf(rf[t], wf[t], acc);
        }
        acc = ctr_warp_sum(acc);
        const float z = acc + (extra ? extra[b] : 0.f) + bs;
        const float p = 1.f / (1.f + expf(-z));
        if (lane == 0) {
            if (logit) logit[b] = z;
            if (score) score[b] = p;
        }
        if (label) {
            const float y = label[b];
            // d/dz [ -y log(p+e) - (1-y) log(1-p+e) ] = (-y/(p+e) + (1-y)/(1-p+e)) * p(1-p), mean over the global batch
            float l_, d_;
            head_loss(p, y, flags, LOG_LOSS_EPS, l_, d_);
            lsum += l_;
            const float dl = d_ * inv_bg;
            if (lane == 0 && dlogit) dlogit[b] = dl;
            if (dz_last) {
                // gradient through the last ReLU: dZ_last[b,j] = dlogit[b] * w_deep[j] * (deep[b,j] > 0)
                float* o = dz_last + b * lddz;
#pragma unroll
                for (int t = 0; t < HEAD_MAXT; ++t) {
                    const int j = lane + 32 * t;
                    if (j < H) {
                        const float g = (!gate || rd[t] > 0.f) ? dl * wd[t] : 0.f;
                        o[j] = round_tf32 ? ctr_round_tf32(g) : g;
                    }
                }
            }
            if (colpart) {
#pragma unroll
                for (int t = 0; t < HEAD_MAXT; ++t) gd[t] = fmaf(dl, rd[t], gd[t]);
#pragma unroll
                for (int t = 0; t < HEAD_MAXF; ++t) gf[t] = fmaf(dl, rf[t], gf[t]);
                gb += dl;
            }
        }
    }
    if (label && colpart) {
        float* mine = sm + w * ncol;
#pragma unroll
        for (int t = 0; t < HEAD_MAXT; ++t) {
            const int j = lane + 32 * t;
            if (j < H) mine[j] = gd[t];
        }
#pragma unroll
        for (int t = 0; t < HEAD_MAXF; ++t) {
            const int j = lane + 32 * t;
            if (j < n_fm) mine[H + j] = gf[t];
        }
        if (lane == 0) mine[H + n_fm] = gb;
        __syncthreads();
        for (int c = threadIdx.x; c < ncol; c += blockDim.x) {
            float a = 0.f;
#pragma unroll
            for (int ww = 0; ww < HEAD_WARPS; ++ww) a += sm[ww * ncol + c];
            colpart[(int64_t) blockIdx.x * ncol + c] = a;
        }
        __syncthreads();
    }
    if (label && loss_part) {
        if (lane == 0) sm[w] = lsum;
        __syncthreads();
        if (threadIdx.x == 0) {
            float a = 0.f;
#pragma unroll
            for (int i = 0; i < HEAD_WARPS; ++i) a += sm[i];
            loss_part[blockIdx.x] = a;
        }
    }
}


// ---- 128-bit variant (H % 4 == 0, 16-byte aligned rows): the kernel above moved 190 MB in 71 us (ncu r01l: 512 blocks on
// 296 slots = 1.73 waves, 13 scalar loads per lane and sample, nothing in flight while a sample is reduced).  Here a
// lane owns float4 column groups (4*lane + 128*t), warps walk the batch with a grid-wide stride (one balanced wave),
// and the NEXT sample's row is already in flight while the current one is reduced and written back.
constexpr int HV_NT = 4;                           // float4 groups per lane of a deep row (H <= 512)
constexpr int HV_WARPS = 8;

__device__ __forceinline__ float f4_dot(const float4 a, const float4 b, float acc) {
    acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc);
    acc = fmaf(a.z, b.z, acc); return fmaf(a.w, b.w, acc);
}

__global__ void __launch_bounds__(256, 2)
head_fwd_vec_kernel(int B, float inv_bg, const float* __restrict__ fm, int64_t fm_ld, int n_fm, const float* __restrict__ w_fm,
                    const float* __restrict__ deep, int64_t deep_ld, int H, const float* __restrict__ w_deep,
                    const float* __restrict__ extra, const float* __restrict__ bias, const float* __restrict__ label,
                    float* __restrict__ logit, float* __restrict__ score, float* __restrict__ dlogit,
                    float* __restrict__ dz_last, int64_t lddz, int round_tf32, float* __restrict__ loss_part,
                    float* __restrict__ colpart, int flags) {
    const bool gate = (flags & CTR_HEAD_NO_RELU_GATE) == 0;
    extern __shared__ float sm[];                  // [HV_WARPS][H + n_fm + 1] when colpart, else [HV_WARPS]
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int ncol = H + n_fm + 1;
    const bool fm_lane = lane * 4 < n_fm;
    float4 wd[HV_NT], gd[HV_NT], cur[HV_NT], nxt[HV_NT];
    float4 wf = make_float4(0.f, 0.f, 0.f, 0.f), gf = wf, curf = wf, nxtf = wf;
#pragma unroll
    for (int t = 0; t < HV_NT; ++t) {
        const int j = 4 * lane + 128 * t;
        // w_deep is a slice of the head weight (offset n_fm floats): not 16-byte aligned in general
        wd[t] = j < H ? make_float4(w_deep[j], w_deep[j + 1], w_deep[j + 2], w_deep[j + 3]) : make_float4(0.f, 0.f, 0.f, 0.f);
        gd[t] = make_float4(0.f, 0.f, 0.f, 0.f);
        cur[t] = nxt[t] = gd[t];
    }
    if (fm_lane) {
        const int j = 4 * lane;                    // w_fm holds n_fm floats: guard the tail of the last group
        wf.x = w_fm[j];
        wf.y = j + 1 < n_fm ? w_fm[j + 1] : 0.f;
        wf.z = j + 2 < n_fm ? w_fm[j + 2] : 0.f;
        wf.w = j + 3 < n_fm ? w_fm[j + 3] : 0.f;
    }
    auto load_row = [&](int64_t b, float4* rd, float4& rf) {
        const float* r = deep + b * deep_ld;
#pragma unroll
        for (int t = 0; t < HV_NT; ++t) {
            const int j = 4 * lane + 128 * t;
            if (j < H) rd[t] = __ldcs(reinterpret_cast<const float4*>(r + j));
        }
        if (fm_lane) {
            rf = __ldcs(reinterpret_cast<const float4*>(fm + b * fm_ld + 4 * lane));
            const int j = 4 * lane;                // pad columns of the fm row are not part of the model
            if (j + 1 >= n_fm) rf.y = 0.f;
            if (j + 2 >= n_fm) rf.z = 0.f;
            if (j + 3 >= n_fm) rf.w = 0.f;
        }
    };
    float gb = 0.f, lsum = 0.f;
    const float bs = bias ? bias[0] : 0.f;
    const int64_t GW = (int64_t) gridDim.x * HV_WARPS;
    int64_t b = (int64_t) blockIdx.x * HV_WARPS + w;
    if (b < B) load_row(b, cur, curf);
    for (; b < B; b += GW) {                       // uniform across the warp
        const int64_t bn = b + GW;
        if (bn < B) load_row(bn, nxt, nxtf);
        float acc = 0.f;
#pragma unroll
        for (int t = 0; t < HV_NT; ++t) acc = f4_dot(cur[t], wd[t], acc);
        acc = f4_dot(curf, wf, acc);
        acc = ctr_warp_sum(acc);
        const float z = acc + (extra ? extra[b] : 0.f) + bs;
        const float p = 1.f / (1.f + expf(-z));
        if (lane == 0) {
            if (logit) logit[b] = z;
            if (score) score[b] = p;
        }
        if (label) {
            const float y = label[b];
            float l_, d_;
            head_loss(p, y, flags, LOG_LOSS_EPS, l_, d_);
            lsum += l_;
            const float dl = d_ * inv_bg;
            if (lane == 0 && dlogit) dlogit[b] = dl;
            if (dz_last) {
                float* o = dz_last + b * lddz;
#pragma unroll
                for (int t = 0; t < HV_NT; ++t) {
                    const int j = 4 * lane + 128 * t;
                    if (j < H) {
                        float4 g;
                        g.x = (!gate || cur[t].x > 0.f) ? dl * wd[t].x : 0.f;
                        g.y = (!gate || cur[t].y > 0.f) ? dl * wd[t].y : 0.f;
                        g.z = (!gate || cur[t].z > 0.f) ? dl * wd[t].z : 0.f;
                        g.w = (!gate || cur[t].w > 0.f) ? dl * wd[t].w : 0.f;
                        if (round_tf32) g = make_float4(ctr_round_tf32(g.x), ctr_round_tf32(g.y), ctr_round_tf32(g.z), ctr_round_tf32(g.w));
                        *reinterpret_cast<float4*>(o + j) = g;
                    }
                }
            }
            if (colpart) {
#pragma unroll
                for (int t = 0; t < HV_NT; ++t) {
                    gd[t].x = fmaf(dl, cur[t].x, gd[t].x); gd[t].y = fmaf(dl, cur[t].y, gd[t].y);
                    gd[t].z = fmaf(dl, cur[t].z, gd[t].z); gd[t].w = fmaf(dl, cur[t].w, gd[t].w);
                }
                gf.x = fmaf(dl, curf.x, gf.x); gf.y = fmaf(dl, curf.y, gf.y);
                gf.z = fmaf(dl, curf.z, gf.z); gf.w = fmaf(dl, curf.w, gf.w);
                gb += dl;
            }
        }
#pragma unroll
        for (int t = 0; t < HV_NT; ++t) cur[t] = nxt[t];
        curf = nxtf;
    }
    if (label && colpart) {
        float* mine = sm + w * ncol;
#pragma unroll
        for (int t = 0; t < HV_NT; ++t) {
            const int j = 4 * lane + 128 * t;
            if (j < H) { mine[j] = gd[t].x; mine[j + 1] = gd[t].y; mine[j + 2] = gd[t].z; mine[j + 3] = gd[t].w; }
        }
        if (fm_lane) {
            const int j = 4 * lane;
            mine[H + j] = gf.x;
            if (j + 1 < n_fm) mine[H + j + 1] = gf.y;
            if (j + 2 < n_fm) mine[H + j + 2] = gf.z;
            if (j + 3 < n_fm) mine[H + j + 3] = gf.w;
        }
        if (lane == 0) mine[H + n_fm] = gb;
        __syncthreads();
        for (int c = threadIdx.x; c < ncol; c += blockDim.x) {
            float a = 0.f;
#pragma unroll
            for (int ww = 0; ww < HV_WARPS; ++ww) a += sm[ww * ncol + c];
            colpart[(int64_t) blockIdx.x * ncol + c] = a;
        }
        __syncthreads();
    }
    if (label && loss_part) {
        if (lane == 0) sm[w] = lsum;
        __syncthreads();
        if (threadIdx.x == 0) {
            float a = 0.f;
#pragma unroll
            for (int i = 0; i < HV_WARPS; ++i) a += sm[i];
            loss_part[blockIdx.x] = a;
        }
    }
}

// ---- the training step's head, streamlined (ncu r02k: the general kernel above spent 99 us on 175 MB, issue-bound: ~50 M
// warp instructions for 64K rows, half of them guards and address arithmetic for cases the training call never has).
// Same lane <-> column mapping, same order of every sum (so the results are those of head_fwd_vec_kernel bit for bit),
// but: every output is present, NT = ceil(H / 128) is a template parameter and only the last float4 group of a row can be
// partial, the label / wide term of the next row travel with its prefetch, and two row buffers alternate instead of
// being copied.
template <int NT>
struct HeadRow {
    float4 d[NT];
    float4 f;
    float y, e;
};

template <int NT, bool GATE, bool ROUND>
__global__ void __launch_bounds__(256, 2)
head_train_kernel(int B, float inv_bg, const float* __restrict__ fm, int64_t fm_ld, int n_fm, const float* __restrict__ w_fm,
                  const float* __restrict__ deep, int64_t deep_ld, int H, const float* __restrict__ w_deep,
                  const float* __restrict__ extra, const float* __restrict__ bias, const float* __restrict__ label,
                  float* __restrict__ logit, float* __restrict__ score, float* __restrict__ dlogit,
                  float* __restrict__ dz_last, int64_t lddz, float* __restrict__ loss_part, float* __restrict__ colpart,
                  int flags) {
    extern __shared__ float sm[];                  // [HV_WARPS][H + n_fm + 1]
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int ncol = H + n_fm + 1;
    const bool fm_lane = lane * 4 < n_fm;
    const bool last_ok = 4 * lane + 128 * (NT - 1) < H;       // H % 4 == 0 and H > 128 * (NT - 1): earlier groups are whole
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 wd[NT], gd[NT];
    float4 wf = zero4, gf = zero4;
#pragma unroll
    for (int t = 0; t < NT; ++t) {
        const int j = 4 * lane + 128 * t;
        wd[t] = j < H ? make_float4(w_deep[j], w_deep[j + 1], w_deep[j + 2], w_deep[j + 3]) : zero4;
        gd[t] = zero4;
    }
    const int jf = 4 * lane;
    const bool f1 = jf + 1 < n_fm, f2 = jf + 2 < n_fm, f3 = jf + 3 < n_fm;
    if (fm_lane) {
        wf.x = w_fm[jf];
        wf.y = f1 ? w_fm[jf + 1] : 0.f;
        wf.z = f2 ? w_fm[jf + 2] : 0.f;
        wf.w = f3 ? w_fm[jf + 3] : 0.f;
    }
    const float bs = bias ? bias[0] : 0.f;
    float gb = 0.f, lsum = 0.f;

    auto load_row = [&](int64_t b, HeadRow<NT>& r) {
        const float4* p = reinterpret_cast<const float4*>(deep + b * deep_ld) + lane;
#pragma unroll
        for (int t = 0; t < NT - 1; ++t) r.d[t] = __ldcs(p + 32 * t);
        r.d[NT - 1] = last_ok ? __ldcs(p + 32 * (NT - 1)) : zero4;
        r.f = zero4;
        if (fm_lane) {
            r.f = __ldcs(reinterpret_cast<const float4*>(fm + b * fm_ld) + lane);
            if (!f1) r.f.y = 0.f;                  // pad columns of the fm row are not part of the model
            if (!f2) r.f.z = 0.f;
            if (!f3) r.f.w = 0.f;
        }
        r.y = label[b];
        r.e = extra ? extra[b] : 0.f;
    };
    auto process = [&](int64_t b, const HeadRow<NT>& r) {
        float acc = 0.f;
#pragma unroll
        for (int t = 0; t < NT; ++t) acc = f4_dot(r.d[t], wd[t], acc);
        acc = f4_dot(r.f, wf, acc);
        acc = ctr_warp_sum(acc);
        const float z = acc + r.e + bs;
        const float p = 1.f / (1.f + expf(-z));
        float l_, d_;
        head_loss(p, r.y, flags, LOG_LOSS_EPS, l_, d_);
        lsum += l_;
        const float dl = d_ * inv_bg;
        if (lane == 0) {
            logit[b] = z;
            score[b] = p;
            dlogit[b] = dl;
        }
        float4* o = reinterpret_cast<float4*>(dz_last + b * lddz) + lane;
#pragma unroll
        for (int t = 0; t < NT; ++t) {
            float4 g = make_float4(dl * wd[t].x, dl * wd[t].y, dl * wd[t].z, dl * wd[t].w);
            // finite values: the same result as cvt.rna.tf32.f32 in two integer operations
            if (ROUND) g = make_float4(ctr_round_tf32_fast(g.x), ctr_round_tf32_fast(g.y), ctr_round_tf32_fast(g.z), ctr_round_tf32_fast(g.w));
            if (GATE) {
                g.x = r.d[t].x > 0.f ? g.x : 0.f;
                g.y = r.d[t].y > 0.f ? g.y : 0.f;
                g.z = r.d[t].z > 0.f ? g.z : 0.f;
                g.w = r.d[t].w > 0.f ? g.w : 0.f;
            }
            if (t < NT - 1 || last_ok) o[32 * t] = g;
            gd[t].x = fmaf(dl, r.d[t].x, gd[t].x); gd[t].y = fmaf(dl, r.d[t].y, gd[t].y);
            gd[t].z = fmaf(dl, r.d[t].z, gd[t].z); gd[t].w = fmaf(dl, r.d[t].w, gd[t].w);
        }
        gf.x = fmaf(dl, r.f.x, gf.x); gf.y = fmaf(dl, r.f.y, gf.y);
        gf.z = fmaf(dl, r.f.z, gf.z); gf.w = fmaf(dl, r.f.w, gf.w);
        gb += dl;
    };

    const int64_t GW = (int64_t) gridDim.x * HV_WARPS;
    int64_t b = (int64_t) blockIdx.x * HV_WARPS + w;
    HeadRow<NT> ra, rb;
    if (b < B) load_row(b, ra);
    while (b < B) {                                // uniform across the warp
        const int64_t b1 = b + GW;
        if (b1 < B) load_row(b1, rb);
        process(b, ra);
        if (b1 >= B) break;
        const int64_t b2 = b1 + GW;
        if (b2 < B) load_row(b2, ra);
        process(b1, rb);
        b = b2;
    }
    float* mine = sm + w * ncol;
#pragma unroll
    for (int t = 0; t < NT; ++t) {
        const int j = 4 * lane + 128 * t;
        if (j < H) { mine[j] = gd[t].x; mine[j + 1] = gd[t].y; mine[j + 2] = gd[t].z; mine[j + 3] = gd[t].w; }
    }
    if (fm_lane) {
        mine[H + jf] = gf.x;
        if (f1) mine[H + jf + 1] = gf.y;
        if (f2) mine[H + jf + 2] = gf.z;
        if (f3) mine[H + jf + 3] = gf.w;
    }
    if (lane == 0) mine[H + n_fm] = gb;
    __syncthreads();
    for (int c = threadIdx.x; c < ncol; c += blockDim.x) {
        float a = 0.f;
#pragma unroll
        for (int ww = 0; ww < HV_WARPS; ++ww) a += sm[ww * ncol + c];
        colpart[(int64_t) blockIdx.x * ncol + c] = a;
    }
    __syncthreads();
    if (lane == 0) sm[w] = lsum;
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f;
#pragma unroll
        for (int i = 0; i < HV_WARPS; ++i) a += sm[i];
        loss_part[blockIdx.x] = a;
    }
}

template <int NT>
static void launch_head_train(int nblk, size_t smem, cudaStream_t st, bool gate, bool round, int B, float inv_bg, const float* fm,
                              int64_t fm_ld, int n_fm, const float* w_fm, const float* deep, int64_t deep_ld, int H,
                              const float* w_deep, const float* extra, const float* bias, const float* label, float* logit,
                              float* score, float* dlogit, float* dz_last, int64_t lddz, float* loss_part, float* colpart,
                              int flags) {
#define CTR_HEAD_TRAIN(G, R)                                                                                              \
    head_train_kernel<NT, G, R><<<nblk, 256, smem, st>>>(B, inv_bg, fm, fm_ld, n_fm, w_fm, deep, deep_ld, H, w_deep, extra, \
                                                         bias, label, logit, score, dlogit, dz_last, lddz, loss_part,      \
                                                         colpart, flags)
    if (gate && round) CTR_HEAD_TRAIN(true, true);
    else if (gate) CTR_HEAD_TRAIN(true, false);
    else if (round) CTR_HEAD_TRAIN(false, true);
    else CTR_HEAD_TRAIN(false, false);
#undef CTR_HEAD_TRAIN
}

static inline int head_vec_blocks(ctr_handle* h, int B) {
    const int64_t want = ctr_cdiv(B > 0 ? B : 1, HV_WARPS), cap = (int64_t) h->sm_count * 2;
    return (int) (want < cap ? want : cap);
}

// column sums of the per-block partial rows -> the three head gradients (one warp per column, fixed order)
__global__ void __launch_bounds__(256)
head_grad_pass2(const float* __restrict__ colpart, int nblk, int H, int n_fm, float* __restrict__ g_deep,
                float* __restrict__ g_fm, float* __restrict__ g_bias) {
    const int ncol = H + n_fm + 1;
    const int col = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (col >= ncol) return;
    float a = 0.f;
    for (int s = lane; s < nblk; s += 32) a += colpart[(int64_t) s * ncol + col];
    a = ctr_warp_sum(a);
    if (lane == 0) {
        if (col < H) { if (g_deep) g_deep[col] = a; }
        else if (col < H + n_fm) { if (g_fm) g_fm[col - H] = a; }
        else if (g_bias) g_bias[0] = a;
    }
}

__global__ void __launch_bounds__(1024) loss_final_kernel(const float* __restrict__ part, int nparts, float inv_bg, float* __restrict__ out) {
    __shared__ float sm[32];
    float acc = 0.f;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) acc += part[i];
    float r = ctr_block_sum(acc, sm);
    if (threadIdx.x == 0) out[0] = r * inv_bg;
}

__global__ void __launch_bounds__(256)
head_bwd_deep_kernel(int B, int H, const float* __restrict__ dlogit, const float* __restrict__ w_deep,
                     const float* __restrict__ deep, int64_t deep_ld, float* __restrict__ dZ, int64_t lddz) {
    const int64_t total = (int64_t) B * H;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t b = t / H;
        const int j = (int) (t - b * H);
        const float a = deep[b * deep_ld + j];
        dZ[b * lddz + j] = a > 0.f ? dlogit[b] * w_deep[j] : 0.f;
    }
}

// column reduction pass 1: block (x = column tile of 128, y = row slice)
constexpr int CS_ROWS = 256;   // rows per block slice
__global__ void __launch_bounds__(128)
colsum_pass1(const float* __restrict__ X, int64_t ldx, int B, int ncols, const float* __restrict__ w,
             float* __restrict__ ws) {
    const int col = blockIdx.x * 128 + threadIdx.x;
    const int64_t lo = (int64_t) blockIdx.y * CS_ROWS;
    const int64_t hi = lo + CS_ROWS < B ? lo + CS_ROWS : B;
    if (col >= ncols) return;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    int64_t b = lo;
    if (w) {
        for (; b + 4 <= hi; b += 4) {
            a0 = fmaf(w[b], X[b * ldx + col], a0);
            a1 = fmaf(w[b + 1], X[(b + 1) * ldx + col], a1);
            a2 = fmaf(w[b + 2], X[(b + 2) * ldx + col], a2);
            a3 = fmaf(w[b + 3], X[(b + 3) * ldx + col], a3);
        }
        for (; b < hi; ++b) a0 = fmaf(w[b], X[b * ldx + col], a0);
    } else {
        for (; b + 4 <= hi; b += 4) {
            a0 += X[b * ldx + col];
            a1 += X[(b + 1) * ldx + col];
            a2 += X[(b + 2) * ldx + col];
            a3 += X[(b + 3) * ldx + col];
        }
        for (; b < hi; ++b) a0 += X[b * ldx + col];
    }
    ws[(int64_t) blockIdx.y * ncols + col] = (a0 + a1) + (a2 + a3);
}

// one warp per column: lane-strided sum over the slices, then the fixed shuffle tree
__global__ void __launch_bounds__(256) colsum_pass2(const float* __restrict__ ws, int nslices, int ncols, float* __restrict__ out) {
    const int col = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (col >= ncols) return;
    float a = 0.f;
    for (int s = lane; s < nslices; s += 32) a += ws[(int64_t) s * ncols + col];
    a = ctr_warp_sum(a);
    if (lane == 0) out[col] = a;
}

}  // namespace

extern "C" int64_t ctr_head_ws_bytes(ctr_handle* h, int B, int n_fm, int H) {
    int64_t nblk = ctr_cdiv(B > 0 ? B : 1, HEAD_SPB);
    if (h && head_vec_blocks(h, B) > nblk) nblk = head_vec_blocks(h, B);      // either kernel may run
    return ctr_align(nblk * 4, 256) + ctr_align(nblk * (int64_t) (H + n_fm + 1) * 4, 256);
}

extern "C" int ctr_head_fwd(ctr_handle* h, ctr_stream s, int B, int64_t B_global, const float* fm, int64_t fm_ld, int n_fm,
                            const float* w_fm, const float* deep, int64_t deep_ld, int H, const float* w_deep,
                            const float* extra, const float* bias, const float* label, float* logit, float* score,
                            float* dlogit, float* dz_last, int64_t lddz, int round_tf32, float* loss_out,
                            float* g_fm, float* g_deep, float* g_bias, void* ws, int64_t ws_bytes, int flags) {
    CTR_REQUIRE(h, B >= 0 && n_fm >= 0 && H >= 0, "head_fwd: negative sizes");
    CTR_REQUIRE(h, n_fm == 0 || (fm && w_fm), "head_fwd: fm inputs null");
    CTR_REQUIRE(h, H == 0 || (deep && w_deep), "head_fwd: deep inputs null");
    CTR_REQUIRE(h, H <= 32 * HEAD_MAXT && n_fm <= 32 * HEAD_MAXF, "head_fwd: H <= %d and n_fm <= %d supported (got %d, %d)",
                32 * HEAD_MAXT, 32 * HEAD_MAXF, H, n_fm);
    if (B == 0) return CTR_OK;
    const float inv_bg = 1.f / (float) (B_global > 0 ? B_global : (int64_t) B);
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool vec = H > 0 && H % 4 == 0 && H <= 128 * HV_NT && deep_ld % 4 == 0 && al16(deep) &&
                     (!dz_last || (lddz % 4 == 0 && al16(dz_last))) && (n_fm == 0 || (fm_ld % 4 == 0 && al16(fm) && fm_ld >= ((n_fm + 3) & ~3)));
    const int nblk = vec ? head_vec_blocks(h, B) : (int) ctr_cdiv(B, HEAD_SPB);
    const bool want_grads = label && (g_fm || g_deep || g_bias);
    float *part = nullptr, *colpart = nullptr;
    if (label && (loss_out || want_grads)) {
        CTR_REQUIRE(h, ws != nullptr, "head_fwd: ws null");
        if (ws_bytes < ctr_head_ws_bytes(h, B, n_fm, H)) return ctr_fail(h, CTR_E_WORKSPACE, "head_fwd: ws too small");
        part = (float*) ws;
        if (want_grads) colpart = (float*) ((char*) ws + ctr_align((int64_t) nblk * 4, 256));
    }
    const size_t smem = (colpart ? (size_t) HEAD_WARPS * (H + n_fm + 1) : (size_t) HEAD_WARPS) * sizeof(float);
    // the training call of the step: every output wanted, whole float4 groups but possibly the last
    const bool train_fast = vec && label && logit && score && dlogit && dz_last && loss_out && colpart &&
                            getenv("CTR_HEAD_GENERAL") == nullptr;
    if (train_fast) {
        const int nt = (H + 127) / 128;
        const bool gate = (flags & CTR_HEAD_NO_RELU_GATE) == 0, rnd = round_tf32 != 0;
#define CTR_HEAD_NT(N)                                                                                                     \
    launch_head_train<N>(nblk, smem, ctr_st(s), gate, rnd, B, inv_bg, fm, fm_ld, n_fm, w_fm, deep, deep_ld, H, w_deep, extra, \
                         bias, label, logit, score, dlogit, dz_last, lddz, part, colpart, flags)
        if (nt == 1) CTR_HEAD_NT(1);
        else if (nt == 2) CTR_HEAD_NT(2);
        else if (nt == 3) CTR_HEAD_NT(3);
        else CTR_HEAD_NT(4);
#undef CTR_HEAD_NT
    } else if (vec)
        head_fwd_vec_kernel<<<nblk, 256, smem, ctr_st(s)>>>(B, inv_bg, fm, fm_ld, n_fm, w_fm, deep, deep_ld, H, w_deep, extra,
                                                            bias, label, logit, score, dlogit, dz_last, lddz, round_tf32,
                                                            loss_out ? part : nullptr, colpart, flags);
    else
        head_fwd_kernel<<<nblk, 256, smem, ctr_st(s)>>>(B, inv_bg, fm, fm_ld, n_fm, w_fm, deep, deep_ld, H, w_deep, extra, bias,
                                                        label, logit, score, dlogit, dz_last, lddz, round_tf32,
                                                        loss_out ? part : nullptr, colpart, flags);
    CTR_CHECK_LAUNCH(h, "head_fwd");
    if (label && loss_out) {
        loss_final_kernel<<<1, 1024, 0, ctr_st(s)>>>(part, nblk, inv_bg, loss_out);
        CTR_CHECK_LAUNCH(h, "loss_final");
    }
    if (colpart) {
        const int ncol = H + n_fm + 1;
        head_grad_pass2<<<(unsigned) ctr_cdiv((int64_t) ncol * 32, 256), 256, 0, ctr_st(s)>>>(colpart, nblk, H, n_fm, g_deep,
                                                                                            g_fm, g_bias);
        CTR_CHECK_LAUNCH(h, "head_grad_pass2");
    }
    return CTR_OK;
}

extern "C" int ctr_head_bwd_deep(ctr_handle* h, ctr_stream s, int B, int H, const float* dlogit, const float* w_deep,
                                 const float* deep, int64_t deep_ld, float* dZ, int64_t lddz) {
    CTR_REQUIRE(h, dlogit && w_deep && deep && dZ && B >= 0 && H > 0, "head_bwd_deep: bad args");
    if (B == 0) return CTR_OK;
    int64_t blocks = ctr_cdiv((int64_t) B * H, 256);
    int64_t cap = (int64_t) h->sm_count * 16;
    if (blocks > cap) blocks = cap;
    head_bwd_deep_kernel<<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(B, H, dlogit, w_deep, deep, deep_ld, dZ, lddz);
    CTR_CHECK_LAUNCH(h, "head_bwd_deep");
    return CTR_OK;
}

extern "C" int64_t ctr_colsum_ws_bytes(ctr_handle*, int B, int ncols) {
    return ctr_cdiv(B > 0 ? B : 1, CS_ROWS) * (int64_t) ncols * 4 + 256;
}

extern "C" int ctr_colsum(ctr_handle* h, ctr_stream s, const float* X, int64_t ldx, int B, int ncols, const float* w,
                          float* out, void* ws, int64_t ws_bytes) {
    CTR_REQUIRE(h, X && out && ws && B >= 0 && ncols > 0 && ldx >= ncols, "colsum: bad args");
    if (ws_bytes < ctr_colsum_ws_bytes(h, B, ncols)) return ctr_fail(h, CTR_E_WORKSPACE, "colsum: ws too small");
    const int nsl = (int) ctr_cdiv(B > 0 ? B : 1, CS_ROWS);
    dim3 grid((unsigned) ctr_cdiv(ncols, 128), (unsigned) nsl);
    colsum_pass1<<<grid, 128, 0, ctr_st(s)>>>(X, ldx, B, ncols, w, (float*) ws);
    CTR_CHECK_LAUNCH(h, "colsum_pass1");
    colsum_pass2<<<(unsigned) ctr_cdiv((int64_t) ncols * 32, 256), 256, 0, ctr_st(s)>>>((const float*) ws, nsl, ncols, out);
    CTR_CHECK_LAUNCH(h, "colsum_pass2");
    return CTR_OK;
}
