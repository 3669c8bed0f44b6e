// Probe (not part of the library): how fast can the SMs fill shared memory with TMA tile loads of the GEMM's shape,
// from L2 and from HBM?  The persistent TF32 GEMMs of mlp_tcgen05.cu stage 28.8 KB per k-block per CTA (A: 128 rows x
// 128 B, B: 100 rows x 128 B, 128B-swizzled boxes out of row-major fp32 matrices with a 1600-byte pitch) and measure
// ~0.64 us per k-block, i.e. ~6.7 TB/s of fill over 148 SMs, most of it L2 hits.  If that is what L2 -> SM delivers,
// the GEMMs are fill-bound and the levers are fewer bytes staged per flop; if the machine can fill much faster, the
// bound is elsewhere (the shared-memory port, the issue order).  DESIGN.md section 8, item 1.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/probe_tma_fill tools/probe_tma_fill.cu && timeout 120 /tmp/probe_tma_fill
//
// One elected thread per CTA keeps a ring of STAGES boxes in flight and never reads them.  Modes:
//   rows = 4096   : the matrix (6.5 MB) stays in L2: pure L2 -> shared-memory fill
//   rows = 1<<20  : 1.7 GB, streamed from HBM
//   shared = 1    : both CTAs of an SM pair (blockIdx / 2) ask for the SAME box (what B-operand sharing looks like)
// Every wait is bounded: a hardware-rejected configuration reports "timeout" instead of hanging the box.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            std::exit(2);                                                                       \
        }                                                                                       \
    } while (0)

constexpr int MAX_STAGES = 8;
constexpr int PITCH = 400;                    // floats per matrix row (1600 B)
constexpr int BOX_W = 32;                     // floats: one 128-byte swizzle row

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

// each iteration: one box of box_rows x 128 B at (k-block kb, row block rb)
__global__ void __launch_bounds__(64) fill_kernel(const __grid_constant__ CUtensorMap tm, int box_rows, int n_rowblocks,
                                                  int iters, int stages, int shared_pairs, unsigned* fail) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    __shared__ __align__(8) uint64_t full[MAX_STAGES];
    const int stage_bytes = box_rows * BOX_W * 4;
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const int who = shared_pairs ? (int) blockIdx.x / 2 : (int) blockIdx.x;
    const int nwho = shared_pairs ? ((int) gridDim.x + 1) / 2 : (int) gridDim.x;
    const int nkb = (PITCH + BOX_W - 1) / BOX_W;          // 13 k-blocks per row block, like the GEMM main loop
    for (int i = 0; i < iters + stages; ++i) {
        const int s = i % stages;
        if (i >= stages) {
            if (!wait_bounded(smem_u32(&full[s]), (uint32_t) ((i / stages) - 1) & 1u)) { atomicAdd(fail, 1u); return; }
        }
        if (i < iters) {
            const int tile = who + (i / nkb) * nwho;      // row blocks walk the matrix like persistent tiles
            const int rb = tile % n_rowblocks, kb = i % nkb;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(stage_bytes) : "memory");
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                ::"r"(smem_u32(smem + (size_t) s * stage_bytes)), "l"(&tm), "r"(smem_u32(&full[s])), "r"(kb * BOX_W),
                  "r"(rb * box_rows) : "memory");
        }
    }
}

// variant kernel: `issuers` warps each run their own ring of `stages` boxes (lane 0 issues); mode 0 = 2D box (one
// k-block), mode 1 = 3D box {32, box_rows, 2} over the same matrix viewed as [k-block][row][32] (two k-blocks per op),
// mode 2 = 1D bulk copy of the same byte count out of a contiguous buffer.
__global__ void __launch_bounds__(256) fill_multi_kernel(const __grid_constant__ CUtensorMap tm, const uint8_t* flat,
                                                         long long flat_bytes, int box_rows, int n_rowblocks, int iters,
                                                         int stages, int issuers, int mode, unsigned* fail) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    __shared__ __align__(8) uint64_t full[32];
    const int op_bytes = box_rows * BOX_W * 4 * (mode == 1 ? 2 : 1);
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages * (issuers < 0 ? -issuers : issuers); ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // issuers > 0: that many warps (lane 0 of each); issuers < 0: that many LANES of warp 0
    const bool by_lane = issuers < 0;
    if (by_lane) issuers = -issuers;
    const int w = by_lane ? (int) threadIdx.x : (int) (threadIdx.x >> 5);
    if ((!by_lane && (threadIdx.x & 31) != 0) || w >= issuers) return;
    const int nkb = 13;
    const int my_iters = iters / issuers;
    for (int i = 0; i < my_iters + stages; ++i) {
        const int s = w * stages + i % stages;
        if (i >= stages) {
            if (!wait_bounded(smem_u32(&full[s]), (uint32_t) ((i / stages) - 1) & 1u)) { atomicAdd(fail, 1u); return; }
        }
        if (i < my_iters) {
            const int g = i * issuers + w;                 // global op index of this CTA
            const int per_tile = mode == 1 ? (nkb / 2) : nkb;
            const int tile = (int) blockIdx.x + (g / per_tile) * (int) gridDim.x;
            const int rb = tile % n_rowblocks, kb = (g % per_tile) * (mode == 1 ? 2 : 1);
            const uint32_t dst = smem_u32(smem + (size_t) s * op_bytes), bar = smem_u32(&full[s]);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(op_bytes) : "memory");
            if (mode == 0) {
                asm volatile(
                    "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                    ::"r"(dst), "l"(&tm), "r"(bar), "r"(kb * BOX_W), "r"(rb * box_rows) : "memory");
            } else if (mode == 1) {
                asm volatile(
                    "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                    ::"r"(dst), "l"(&tm), "r"(bar), "r"(0), "r"(rb * box_rows), "r"(kb) : "memory");
            } else {
                const long long off = ((long long) tile * nkb + kb) * op_bytes % (flat_bytes - op_bytes);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(dst), "l"(flat + (off & ~15ll)), "r"(op_bytes), "r"(bar) : "memory");
            }
        }
    }
}

// part 3: do the epilogue's global stores slow the TMA loads down, and does it matter how they are issued?
// warp 0: four lanes keep 2D boxes of 228 rows x 128 B (one k-block's A + B bytes) in flight, 13 per "unit";
// warps 1-8: per unit each writes three 32 x 32-float tiles of a [rows, pitch] output (24 tiles ~ the 128 x 200 floats a
// CTA stores per unit), smode 1 = coalesced st.global.v4 (4 rows x 128 B per instruction, like epilogue_rows_fast),
// smode 2 = one TMA tile store per tile out of shared memory.  what & 1: loads on, what & 2: stores on.
__global__ void __launch_bounds__(288) mix_kernel(const __grid_constant__ CUtensorMap tmL, const __grid_constant__ CUtensorMap tmS,
                                                  float* out, int out_pitch, int out_rows, int n_rowblocks, int units, int what,
                                                  int smode, unsigned* fail) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    __shared__ __align__(8) uint64_t full[8];
    constexpr int LANES = 4, STAGES = 4, BOXR = 228;
    const int op_bytes = BOXR * BOX_W * 4;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (w == 0) {
        if (!(what & 1) || lane >= LANES) return;
        const int total = units * 13;
        for (int i = lane; i < total + STAGES; i += LANES) {     // lane l owns stage l
            const int s = lane;
            const int use = i / LANES;
            if (use >= 1) {
                if (!wait_bounded(smem_u32(&full[s]), (uint32_t) (use - 1) & 1u)) { atomicAdd(fail, 1u); return; }
            }
            if (i < total) {
                const int tile = (int) blockIdx.x + (i / 13) * (int) gridDim.x;
                const int rb = tile % n_rowblocks, kb = i % 13;
                const uint32_t bar = smem_u32(&full[s]);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(op_bytes) : "memory");
                asm volatile(
                    "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                    ::"r"(smem_u32(smem + (size_t) s * op_bytes)), "l"(&tmL), "r"(bar), "r"(kb * BOX_W), "r"(rb * BOXR) : "memory");
            }
        }
        return;
    }
    if (!(what & 2)) return;
    const int ew = w - 1;                                   // 0..7
    float* tile_s = reinterpret_cast<float*>(smem + (size_t) STAGES * op_bytes + (size_t) ew * 4096);
    for (int i = lane; i < 1024; i += 32) tile_s[i] = (float) i;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    const int rsub = lane >> 3, cc = (lane & 7) * 4;
    for (int u = 0; u < units; ++u) {
        const int tile = (int) blockIdx.x + u * (int) gridDim.x;
        const int row0 = (tile * 128 + (ew & 3) * 32) % (out_rows - 32);
        for (int t = 0; t < 3; ++t) {
            const int col0 = ((ew >> 2) * 3 + t) * 32 + (u & 1) * 192;          // 6 chunks of 32 columns per half
            if (smode == 1) {
                const float4 v = make_float4(1.f, 2.f, 3.f, (float) u);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<float4*>(out + (size_t) (row0 + rsub + 4 * j) * out_pitch + col0 + cc) = v;
            } else {
                if (lane == 0) {
                    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                                 ::"l"(&tmS), "r"(smem_u32(tile_s)), "r"(col0), "r"(row0) : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                }
                __syncwarp();
            }
        }
    }
    if (smode == 2 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
    const bool only3 = argc > 1 && argv[1][0] == '3';
    CK(cudaSetDevice(0));
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    encode_fn enc = (encode_fn) p;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    unsigned* fail;
    CK(cudaMalloc(&fail, 4));
    CK(cudaMemset(fail, 0, 4));
    CK(cudaFuncSetAttribute(fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    std::printf("%d SMs.  box = rows x 128 B out of a [rows, 400] fp32 matrix; one issuing thread per CTA\n", sms);
    std::printf("%9s %5s %7s %7s %7s %9s %10s %9s\n", "rows", "box", "stages", "cta/sm", "shared", "us", "TB/s", "GB/s/SM");
    const long long rows_list[2] = {4096, 1ll << 20};
    for (int ri = 0; ri < (only3 ? 0 : 2); ++ri) {
        const long long rows = rows_list[ri];
        float* mat;
        CK(cudaMalloc(&mat, (size_t) rows * PITCH * 4));
        CK(cudaMemset(mat, 0, (size_t) rows * PITCH * 4));
        const int boxes[3] = {128, 228, 64};
        for (int bi = 0; bi < 3; ++bi) {
            const int box_rows = boxes[bi];
            CUtensorMap tm;
            cuuint64_t dims[2] = {(cuuint64_t) PITCH, (cuuint64_t) rows};
            cuuint64_t strides[1] = {(cuuint64_t) PITCH * 4};
            cuuint32_t box[2] = {(cuuint32_t) BOX_W, (cuuint32_t) box_rows};
            cuuint32_t estr[2] = {1, 1};
            CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, mat, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { std::printf("encode failed %d\n", (int) r); continue; }
            const int stage_bytes = box_rows * BOX_W * 4;
            for (int stages = 2; stages <= 6; stages += 2) {
                for (int per_sm = 1; per_sm <= 2; ++per_sm) {
                    for (int shared = 0; shared <= 1; ++shared) {
                        const size_t smem = (size_t) stages * stage_bytes + 1024;
                        if (smem * per_sm > 220 * 1024) continue;
                        const int grid = sms * per_sm;
                        const int iters = 13 * 64;
                        const int n_rowblocks = (int) (rows / box_rows);
                        for (int rep = 0; rep < 3; ++rep) {      // the last repetition is the one printed
                            CK(cudaEventRecord(e0));
                            fill_kernel<<<grid, 64, smem>>>(tm, box_rows, n_rowblocks, iters, stages, shared, fail);
                            CK(cudaEventRecord(e1));
                            CK(cudaEventSynchronize(e1));
                        }
                        CK(cudaGetLastError());
                        float ms = 0;
                        CK(cudaEventElapsedTime(&ms, e0, e1));
                        unsigned f = 0;
                        CK(cudaMemcpy(&f, fail, 4, cudaMemcpyDeviceToHost));
                        if (f) { std::printf("timeout (%u CTAs)\n", f); CK(cudaMemset(fail, 0, 4)); continue; }
                        const double bytes = (double) grid * iters * stage_bytes;
                        std::printf("%9lld %5d %7d %7d %7d %9.1f %10.2f %9.1f\n", rows, box_rows, stages, per_sm, shared, ms * 1e3,
                                    bytes / ms / 1e9, bytes / ms / 1e6 / sms);
                    }
                }
            }
        }
        CK(cudaFree(mat));
    }
    // ---- part 2: what lifts the per-op time?  several issuing warps, a 2048-byte pitch, two k-blocks per op, 1D bulk
    CK(cudaFuncSetAttribute(fill_multi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    std::printf("\npart 2 (4096 rows, L2-resident; 1 CTA per SM): %6s %6s %8s %7s %9s %9s %10s\n", "pitch", "mode", "issuers", "stages",
                "us", "GB/s/SM", "ns/op/CTA");
    for (int pitch = 400; pitch <= (only3 ? 0 : 512); pitch += 112) {
        const long long rows = 4096;
        float* mat;
        CK(cudaMalloc(&mat, (size_t) rows * pitch * 4 + 4096));
        CK(cudaMemset(mat, 0, (size_t) rows * pitch * 4 + 4096));
        for (int mode = 0; mode <= 2; ++mode) {
            const int box_rows = 128;
            CUtensorMap tm;
            CUresult r;
            cuuint32_t estr[3] = {1, 1, 1};
            if (mode == 1) {
                cuuint64_t dims[3] = {(cuuint64_t) BOX_W, (cuuint64_t) rows, (cuuint64_t) 12};
                cuuint64_t strides[2] = {(cuuint64_t) pitch * 4, (cuuint64_t) BOX_W * 4};
                cuuint32_t box[3] = {(cuuint32_t) BOX_W, (cuuint32_t) box_rows, 2};
                r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, mat, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            } else {
                cuuint64_t dims[2] = {(cuuint64_t) pitch, (cuuint64_t) rows};
                cuuint64_t strides[1] = {(cuuint64_t) pitch * 4};
                cuuint32_t box[2] = {(cuuint32_t) BOX_W, (cuuint32_t) box_rows};
                r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, mat, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            }
            if (r != CUDA_SUCCESS) { std::printf("encode failed %d (pitch %d mode %d)\n", (int) r, pitch, mode); continue; }
            const int op_bytes = box_rows * BOX_W * 4 * (mode == 1 ? 2 : 1);
            const int issuer_list[5] = {1, 2, 4, -2, -4};
            for (int ii = 0; ii < 5; ++ii) {
                const int issuers = issuer_list[ii], n_iss = issuers < 0 ? -issuers : issuers;
                for (int stages = 2; stages <= 4; stages += 2) {
                    const size_t smem = (size_t) stages * n_iss * op_bytes + 1024;
                    if (smem > 200 * 1024) continue;
                    const int iters = 12 * 64;
                    for (int rep = 0; rep < 3; ++rep) {
                        CK(cudaEventRecord(e0));
                        fill_multi_kernel<<<sms, 256, smem>>>(tm, reinterpret_cast<const uint8_t*>(mat), rows * pitch * 4, box_rows,
                                                               (int) (rows / box_rows), iters, stages, issuers, mode, fail);
                        CK(cudaEventRecord(e1));
                        CK(cudaEventSynchronize(e1));
                    }
                    CK(cudaGetLastError());
                    float ms = 0;
                    CK(cudaEventElapsedTime(&ms, e0, e1));
                    unsigned f = 0;
                    CK(cudaMemcpy(&f, fail, 4, cudaMemcpyDeviceToHost));
                    if (f) { std::printf("timeout (%u CTAs)\n", f); CK(cudaMemset(fail, 0, 4)); continue; }
                    const double bytes = (double) sms * iters * op_bytes;
                    std::printf("%54d %6d %8d %7d %9.1f %9.1f %10.1f\n", pitch, mode, issuers, stages, ms * 1e3, bytes / ms / 1e6 / sms,
                                ms * 1e6 / iters);
                }
            }
        }
        CK(cudaFree(mat));
    }
    // ---- part 3: loads and stores together
    {
        CK(cudaFuncSetAttribute(mix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        const long long rows = 1ll << 16;
        float *mat, *out;
        CK(cudaMalloc(&mat, (size_t) rows * PITCH * 4));
        CK(cudaMemset(mat, 0, (size_t) rows * PITCH * 4));
        std::printf("\npart 3 (A 64K x 400 from HBM, 148 CTAs, 14 units each): what 1 = loads, 2 = stores, 3 = both\n");
        std::printf("%10s %6s %6s %9s\n", "out pitch", "smode", "what", "us");
        for (int out_pitch = 404; out_pitch <= 416; out_pitch += 12) {
            CK(cudaMalloc(&out, (size_t) rows * out_pitch * 4));
            CUtensorMap tmL, tmS;
            cuuint32_t estr[2] = {1, 1};
            {
                cuuint64_t dims[2] = {(cuuint64_t) PITCH, (cuuint64_t) rows};
                cuuint64_t strides[1] = {(cuuint64_t) PITCH * 4};
                cuuint32_t box[2] = {(cuuint32_t) BOX_W, 228};
                CUresult r = enc(&tmL, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, mat, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (r != CUDA_SUCCESS) { std::printf("encode L failed %d\n", (int) r); return 1; }
            }
            {
                cuuint64_t dims[2] = {(cuuint64_t) 400, (cuuint64_t) rows};
                cuuint64_t strides[1] = {(cuuint64_t) out_pitch * 4};
                cuuint32_t box[2] = {32, 32};
                CUresult r = enc(&tmS, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, out, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                if (r != CUDA_SUCCESS) { std::printf("encode S failed %d\n", (int) r); return 1; }
            }
            const size_t smem = (size_t) 4 * 228 * 128 + 8 * 4096 + 1024;
            for (int smode = 1; smode <= 2; ++smode) {
                for (int what = 1; what <= 3; ++what) {
                    if (what == 1 && smode == 2) continue;
                    for (int rep = 0; rep < 3; ++rep) {
                        CK(cudaEventRecord(e0));
                        mix_kernel<<<sms, 288, smem>>>(tmL, tmS, out, out_pitch, (int) rows, (int) (rows / 228), 14, what, smode, fail);
                        CK(cudaEventRecord(e1));
                        CK(cudaEventSynchronize(e1));
                    }
                    CK(cudaGetLastError());
                    float ms = 0;
                    CK(cudaEventElapsedTime(&ms, e0, e1));
                    unsigned f = 0;
                    CK(cudaMemcpy(&f, fail, 4, cudaMemcpyDeviceToHost));
                    if (f) { std::printf("timeout (%u CTAs)\n", f); CK(cudaMemset(fail, 0, 4)); continue; }
                    std::printf("%10d %6d %6d %9.1f\n", out_pitch, smode, what, ms * 1e3);
                }
            }
            CK(cudaFree(out));
        }
        CK(cudaFree(mat));
    }
    return 0;
}
