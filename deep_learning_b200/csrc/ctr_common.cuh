// Shared device/host helpers for libctr_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include "../../include/ctr_b200.h"

struct ctr_handle {
    int device;
    int sm_count;
    int cc_major, cc_minor;
    unsigned long long launches;   // kernels launched through this handle
    char err[512];
    float* reduce_scratch;         // device address of ctr_api.cu's static scratch ON THIS HANDLE'S DEVICE (ctr_reduce_static)
    int reduce_scratch_rc;         // cudaError_t of the cudaGetSymbolAddress that produced it
};

#define CTR_PAD_KEY 0xFFFFFFFFu

static inline int ctr_fail(ctr_handle* h, int code, const char* fmt, ...) {
    if (h) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(h->err, sizeof(h->err), fmt, ap);
        va_end(ap);
    }
    return code;
}

#define CTR_REQUIRE(h, cond, ...)                                  \
    do {                                                           \
        if (!(cond)) return ctr_fail((h), CTR_E_INVALID, __VA_ARGS__); \
    } while (0)

// checks the launch (not the execution): keeps every API call asynchronous
#define CTR_CHECK_LAUNCH(h, what)                                                        \
    do {                                                                                 \
        cudaError_t _e = cudaGetLastError();                                             \
        if (_e != cudaSuccess)                                                           \
            return ctr_fail((h), CTR_E_CUDA, "%s: %s", (what), cudaGetErrorString(_e));  \
        (h)->launches += 1;                                                              \
    } while (0)

#define CTR_CUDA(h, call)                                                                \
    do {                                                                                 \
        cudaError_t _e = (call);                                                         \
        if (_e != cudaSuccess)                                                           \
            return ctr_fail((h), CTR_E_CUDA, "%s: %s", #call, cudaGetErrorString(_e));   \
    } while (0)

static inline cudaStream_t ctr_st(ctr_stream s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int64_t ctr_cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t ctr_align(int64_t a, int64_t b) { return ctr_cdiv(a, b) * b; }

// ---------------------------------------------------------------------------- device side
#ifdef __CUDACC__

// an id outside its table: remember the field (status word CTR_STATUS_BAD_ID; only ever set on the error path)
__device__ __forceinline__ void ctr_flag_bad_id(uint32_t* status, int field) {
    if (status) atomicMax(status + CTR_STATUS_BAD_ID, 1u + (uint32_t) field);
}
// optimizer kernels: true while the engine's status words say the step must not move the weights
__device__ __forceinline__ bool ctr_guard_tripped(const uint32_t* guard) {
    return guard && (guard[CTR_STATUS_STEP_BARRIER] | guard[CTR_STATUS_PLAN_BARRIER] | guard[CTR_STATUS_BAD_ID] |
                     guard[CTR_STATUS_EXCHANGE]) != 0u;
}

__device__ __forceinline__ int64_t ctr_load_id(const void* ids, int id_bytes, int64_t i) {
    return id_bytes == 8 ? reinterpret_cast<const int64_t*>(ids)[i]
                         : (int64_t) reinterpret_cast<const int32_t*>(ids)[i];
}

// streaming 128-bit load that does not pollute L1 (rows are used once per CTA)
__device__ __forceinline__ float4 ctr_ldg_f4_stream(const float* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float ctr_ldg_f1_stream(const float* p) {
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}

// round-to-nearest fp32 -> tf32 (10-bit mantissa), kept in an fp32 container.  The tensor core TRUNCATES
// fp32 operands to tf32; producers on the TF32 path pre-round so the operand error is unbiased (2^-11).
__device__ __forceinline__ float ctr_round_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

// the same rounding for finite values in two integer ops (ptxas expands the cvt into ~4 with inf/nan handling);
// inf stays inf, canonical NaNs stay NaN
__device__ __forceinline__ float ctr_round_tf32_fast(float x) {
    return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}

__device__ __forceinline__ float ctr_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// deterministic block sum (fixed tree); result valid in thread 0. blockDim.x multiple of 32, <= 1024
__device__ __forceinline__ float ctr_block_sum(float v, float* smem32) {
    v = ctr_warp_sum(v);
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) smem32[w] = v;
    __syncthreads();
    float r = 0.f;
    if (w == 0) {
        int nw = (blockDim.x + 31) >> 5;
        r = lane < nw ? smem32[lane] : 0.f;
        r = ctr_warp_sum(r);
    }
    __syncthreads();
    return r;
}

#endif  // __CUDACC__
