// Probe for the next GPU visit (not part of the library): can GEMM-0 read its A operand straight from the embedding
// table with TMA tile::gather4, and at what rate?  DESIGN.md section 8, item 1.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/probe_gather4 tools/probe_gather4.cu && timeout 120 /tmp/probe_gather4
//
// For box heights 1 and 4 and swizzle NONE / 64B it (1) checks what one gather4 of 4 arbitrary rows x 16 floats leaves in
// shared memory against the table, (2) times a ring of such operations issued by one thread per CTA, many CTAs per SM,
// over random rows of a 1M x 24-float table, and prints ns per operation per SM and GB/s.  Every wait is bounded: a
// configuration the hardware rejects reports "timeout" instead of hanging the box.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            std::exit(2);                                                                       \
        }                                                                                       \
    } while (0)

constexpr int STAGES = 8;
constexpr int WIDTH = 16;                     // floats per row taken (one field's embedding, 64 B)
constexpr int STAGE_BYTES = 4 * WIDTH * 4;    // 4 rows
constexpr int TABLE_STRIDE = 24;              // floats (the library's row stride for k = 16 + first-order weight)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ bool wait_bounded(uint32_t bar, uint32_t parity) {
    for (int spin = 0; spin < (1 << 22); ++spin) {
        uint32_t ok;
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return true;
    }
    return false;
}

__device__ __forceinline__ void gather4(uint32_t dst, const CUtensorMap* tm, int col, int r0, int r1, int r2, int r3, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
                 " [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                 :: "r"(dst), "l"(tm), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(bar) : "memory");
}

// one thread per CTA keeps STAGES operations in flight; groups[] holds 4 row numbers per operation
__global__ void __launch_bounds__(32) ring_kernel(const __grid_constant__ CUtensorMap tm, const int* __restrict__ groups,
                                                  int groups_per_cta, float* __restrict__ dump, int* __restrict__ status) {
    __shared__ __align__(1024) unsigned char stage[STAGES * STAGE_BYTES];
    __shared__ __align__(8) uint64_t bar[STAGES];
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const int* my = groups + (size_t) blockIdx.x * groups_per_cta * 4;
        for (int g = 0; g < groups_per_cta + STAGES; ++g) {
            const int s = g % STAGES;
            if (g >= STAGES) {
                if (!wait_bounded(smem_u32(&bar[s]), (uint32_t) ((g / STAGES - 1) & 1))) {
                    atomicExch(status, 1);
                    return;
                }
            }
            if (g < groups_per_cta) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(&bar[s])), "r"(STAGE_BYTES) : "memory");
                gather4(smem_u32(stage + s * STAGE_BYTES), &tm, 0, my[4 * g], my[4 * g + 1], my[4 * g + 2], my[4 * g + 3], smem_u32(&bar[s]));
            }
        }
    }
    __syncthreads();
    if (dump && blockIdx.x == 0)          // what the LAST operation of every stage left behind (groups_per_cta <= STAGES when dumping)
        for (int i = threadIdx.x; i < STAGES * STAGE_BYTES / 4; i += blockDim.x) dump[i] = reinterpret_cast<float*>(stage)[i];
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    const int64_t R = 1 << 20;
    int dev = 0, sms = 0;
    CK(cudaSetDevice(dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    if (!fn) { std::printf("no cuTensorMapEncodeTiled\n"); return 2; }
    EncodeTiled encode = (EncodeTiled) fn;

    std::vector<float> h_table((size_t) R * TABLE_STRIDE);
    for (size_t i = 0; i < h_table.size(); ++i) h_table[i] = (float) (i % 1000003) * 0.5f;
    float* d_table;
    CK(cudaMalloc(&d_table, h_table.size() * 4));
    CK(cudaMemcpy(d_table, h_table.data(), h_table.size() * 4, cudaMemcpyHostToDevice));

    const int ctas_per_sm = 8, groups_per_cta = 512;
    const int grid = sms * ctas_per_sm;
    std::vector<int> h_groups((size_t) grid * groups_per_cta * 4);
    uint64_t x = 88172645463325252ull;
    for (auto& v : h_groups) { x ^= x << 13; x ^= x >> 7; x ^= x << 17; v = (int) (x % R); }
    int* d_groups;
    CK(cudaMalloc(&d_groups, h_groups.size() * 4));
    CK(cudaMemcpy(d_groups, h_groups.data(), h_groups.size() * 4, cudaMemcpyHostToDevice));
    float* d_dump;
    int* d_status;
    CK(cudaMalloc(&d_dump, STAGES * STAGE_BYTES));
    CK(cudaMalloc(&d_status, 4));

    const CUtensorMapSwizzle swz[2] = {CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_SWIZZLE_64B};
    const char* swz_name[2] = {"none", "64B"};
    for (int box_h = 1; box_h <= 4; box_h += 3) {
        for (int si = 0; si < 2; ++si) {
            CUtensorMap tm;
            cuuint64_t gdim[2] = {(cuuint64_t) TABLE_STRIDE, (cuuint64_t) R};
            cuuint64_t gstr[1] = {(cuuint64_t) TABLE_STRIDE * 4};
            cuuint32_t box[2] = {(cuuint32_t) WIDTH, (cuuint32_t) box_h};
            cuuint32_t estr[2] = {1, 1};
            CUresult r = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d_table, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                swz[si], CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            std::printf("box {%d,%d} swizzle %-4s: encode rc %d", WIDTH, box_h, swz_name[si], (int) r);
            if (r != CUDA_SUCCESS) { std::printf("\n"); continue; }
            // (1) content of one operation
            CK(cudaMemset(d_status, 0, 4));
            CK(cudaMemset(d_dump, 0xff, STAGES * STAGE_BYTES));
            ring_kernel<<<1, 32>>>(tm, d_groups, 1, d_dump, d_status);
            cudaError_t e = cudaDeviceSynchronize();
            int st = 0;
            if (e != cudaSuccess) { std::printf("  launch error: %s\n", cudaGetErrorString(e)); return 3; }
            CK(cudaMemcpy(&st, d_status, 4, cudaMemcpyDeviceToHost));
            if (st) { std::printf("  timeout (operation never completed)\n"); continue; }
            std::vector<float> got(STAGE_BYTES / 4);
            CK(cudaMemcpy(got.data(), d_dump, STAGE_BYTES, cudaMemcpyDeviceToHost));
            int linear_ok = 1;
            for (int i = 0; i < 4; ++i)
                for (int j = 0; j < WIDTH; ++j)
                    if (got[i * WIDTH + j] != h_table[(size_t) h_groups[i] * TABLE_STRIDE + j]) linear_ok = 0;
            // with a swizzle the 16-byte chunks of a row are permuted inside the 64-byte line: check as multisets per row
            int rows_ok = 1;
            for (int i = 0; i < 4; ++i) {
                for (int j = 0; j < WIDTH; ++j) {
                    const float want = h_table[(size_t) h_groups[i] * TABLE_STRIDE + j];
                    int found = 0;
                    for (int t = 0; t < 4 * WIDTH; ++t) found |= got[t] == want;
                    rows_ok &= found;
                }
            }
            std::printf("  rows present %d, row-major unswizzled %d", rows_ok, linear_ok);
            // (2) rate
            cudaEvent_t a, b;
            CK(cudaEventCreate(&a));
            CK(cudaEventCreate(&b));
            ring_kernel<<<grid, 32>>>(tm, d_groups, groups_per_cta, nullptr, d_status);          // warm-up
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(a));
            ring_kernel<<<grid, 32>>>(tm, d_groups, groups_per_cta, nullptr, d_status);
            CK(cudaEventRecord(b));
            CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(&st, d_status, 4, cudaMemcpyDeviceToHost));
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, a, b));
            const double ops = (double) grid * groups_per_cta;
            std::printf("  | %d CTAs x %d ops: %.1f us, %.1f ns per op per SM, %.0f GB/s%s\n", grid, groups_per_cta, ms * 1e3,
                        ms * 1e6 / (ops / sms), ops * STAGE_BYTES / (ms * 1e-3) / 1e9, st ? "  (TIMEOUT in the timed run)" : "");
        }
    }
    // reference point: the library's gather kernel moves 64K x 26 rows in about 47 us = 2 880 gather4-sized operations per SM
    std::printf("need: <= ~16 ns per operation per SM to match the fused gather kernel (47 us for 64K x 26 rows / 4 on %d SMs)\n", sms);
    return 0;
}
