// GPU AUC: exact tie-aware rank statistic, equal to sklearn.metrics.roc_auc_score up to the final
// fp64 division (models/deepfm.py:231,253).  Scores are mapped to order-preserving uint32 keys,
// radix-sorted ascending; for every group of tied scores
//     2*U += npos_g * (2 * negatives_before_g + nneg_g)
// in 64-bit integers (atomicAdd on integers is order-independent, hence deterministic).
#include "ctr_common.cuh"

int ctr_scan_pos_labels(ctr_handle* h, cudaStream_t st, const float* labels, const uint32_t* perm, int64_t n,
                        uint32_t* out, uint32_t* blk_ws);   // sort_scan.cu

namespace {

__global__ void auc_keys_kernel(const float* __restrict__ scores, int64_t n, uint32_t* __restrict__ keys,
                                uint32_t* __restrict__ idx) {
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = scores[i];
    if (s == 0.f) s = 0.f;   // -0.0 == +0.0 must tie
    uint32_t u = __float_as_uint(s);
    keys[i] = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    idx[i] = (uint32_t) i;
}

__global__ void auc_groups_kernel(const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ n_groups,
                                  const uint32_t* __restrict__ cumpos, unsigned long long* __restrict__ acc) {
    const int64_t G = *n_groups;
    unsigned long long local = 0;
    for (int64_t g = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; g < G; g += (int64_t) gridDim.x * blockDim.x) {
        const uint32_t s = seg_start[g], e = seg_start[g + 1];
        const unsigned long long pos_before = cumpos[s];
        const unsigned long long npos = cumpos[e] - cumpos[s];
        const unsigned long long nneg = (unsigned long long) (e - s) - npos;
        const unsigned long long neg_before = (unsigned long long) s - pos_before;
        local += npos * (2ull * neg_before + nneg);
    }
    // warp-aggregate then one integer atomic per warp
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(acc, local);
}

__global__ void auc_final_kernel(const unsigned long long* __restrict__ acc, const uint32_t* __restrict__ cumpos, int64_t n,
                                 double* __restrict__ out) {
    const double P = (double) cumpos[n], N = (double) n - (double) cumpos[n];
    out[0] = (P > 0 && N > 0) ? (double) acc[0] / (2.0 * P * N) : nan("");
}

}  // namespace

extern "C" int64_t ctr_auc_ws_bytes(ctr_handle* h, int64_t n) {
    // keys, idx, uniq, seg_start(+1), uidx, cumpos(+1), scan blocks, acc + sort ws
    return ctr_align(n * 4, 256) * 3 + ctr_align((n + 1) * 4, 256) * 2 + ctr_align((n / 2048 + 4) * 4, 256) + 512 +
           ctr_sort_ws_bytes(h, n);
}

extern "C" int ctr_auc(ctr_handle* h, ctr_stream s, const float* scores, const float* labels, int64_t n, double* auc_out,
                       void* ws, int64_t ws_bytes) {
    CTR_REQUIRE(h, scores && labels && auc_out && ws && n > 0 && n < ((int64_t) 1 << 31), "auc: bad args");
    if (ws_bytes < ctr_auc_ws_bytes(h, n)) return ctr_fail(h, CTR_E_WORKSPACE, "auc: ws too small");
    cudaStream_t st = ctr_st(s);
    char* p = (char*) ws;
    uint32_t* keys = (uint32_t*) p; p += ctr_align(n * 4, 256);
    uint32_t* idx = (uint32_t*) p; p += ctr_align(n * 4, 256);
    uint32_t* uidx = (uint32_t*) p; p += ctr_align(n * 4, 256);
    uint32_t* seg = (uint32_t*) p; p += ctr_align((n + 1) * 4, 256);
    uint32_t* cum = (uint32_t*) p; p += ctr_align((n + 1) * 4, 256);
    uint32_t* blk = (uint32_t*) p; p += ctr_align((n / 2048 + 4) * 4, 256);
    unsigned long long* acc = (unsigned long long*) p; p += 256;
    uint32_t* ngroups = (uint32_t*) p; p += 256;
    void* sort_ws = p;
    int64_t sort_bytes = ws_bytes - (p - (char*) ws);
    auc_keys_kernel<<<(unsigned) ctr_cdiv(n, 256), 256, 0, st>>>(scores, n, keys, idx);
    CTR_CHECK_LAUNCH(h, "auc_keys");
    int rc = ctr_sort_pairs(h, s, keys, idx, n, 32, sort_ws, sort_bytes);
    if (rc) return rc;
    // uniq output is not needed: alias it onto the (now free) sort workspace head
    rc = ctr_unique_segments(h, s, keys, n, (uint32_t*) sort_ws, seg, uidx, ngroups, blk, (n / 2048 + 4) * 4, nullptr);
    if (rc) return rc;
    rc = ctr_scan_pos_labels(h, st, labels, idx, n, cum, blk);
    if (rc) return rc;
    CTR_CUDA(h, cudaMemsetAsync(acc, 0, 8, st));
    int64_t blocks = ctr_cdiv(n, 256);
    if (blocks > (int64_t) h->sm_count * 8) blocks = (int64_t) h->sm_count * 8;
    auc_groups_kernel<<<(unsigned) blocks, 256, 0, st>>>(seg, ngroups, cum, acc);
    auc_final_kernel<<<1, 1, 0, st>>>(acc, cum, n, auc_out);
    h->launches += 1;
    CTR_CHECK_LAUNCH(h, "auc");
    return CTR_OK;
}
