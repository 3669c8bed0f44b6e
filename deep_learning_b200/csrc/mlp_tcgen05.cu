// K3 (TF32 path): hand-written tcgen05 / TMA / TMEM GEMMs for the MLP tower (sm_100a).
//
//   tn kernel  C[M,N] = A[M,R] * B[N,R]^T      both operands reduction-major ("K-major"):
//              forward  (A = X,  B = W^T copy) with bias + ReLU (+dropout) epilogue
//              dgrad    (A = dZ, B = W)        with ReLU/dropout-mask epilogue
//   nt kernel  C[M,N] = A[R,M]^T * B[R,N]      both operands MN-major, split over the reduction:
//              wgrad    (A = X, B = dZ, R = batch)
//
// One CTA owns a 128-row output panel across the WHOLE N extent (N <= 512 fp32 accumulator columns =
// all of TMEM), so the A panel is streamed from HBM exactly once.  Warp roles: warps 0-7 epilogue
// (TMEM lane quarter = warp id & 3, alternate 32-column chunks), warp 8 = TMA producer (one elected
// lane), warp 9 = MMA issuer (one elected lane).  Operands are staged by TMA into 128B-swizzled shared memory through a 3-stage
// mbarrier ring; accumulators live in TMEM and are read back with tcgen05.ld for the fused epilogue.
#include <cuda.h>
#include <stdlib.h>
#include "ctr_common.cuh"

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_R = 32;            // reduction elements per stage: 32 fp32 = one 128B swizzle row
constexpr int UMMA_K = 8;              // tf32: 32 bytes of reduction per instruction
constexpr int MAX_N = 512;             // fp32 accumulator columns in TMEM
// warp roles: warps [0, EW) epilogue (EW = 8 or 16), warp EW = TMA producer, warp EW + 1 = MMA issuer
constexpr int nthreads(int ew) { return (ew + 2) * 32; }
constexpr int SBIAS_BYTES = 2048;          // bias of the CTA's columns staged in shared memory (N <= 512)
constexpr int SMEM_LIMIT = 227 * 1024;

// ------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// bounded wait: a protocol error traps (CUDA error) instead of hanging the GPU box
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (!done && spin > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem desc] * B[smem desc], kind::tf32, issued by one thread
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on an mbarrier once every MMA issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// ---- CTA-pair (cta_group::2) variants: two CTAs of a cluster on the two SMs of a TPC run ONE M=256 MMA; each CTA
// stages its own 128 rows of A and HALF of the B tile, so the shared-memory fill per SM per k-block drops from
// (128 + N) to (128 + N/2) rows.  Only the leader (cluster rank 0) issues MMAs; both CTAs issue TMA, all transaction
// bytes are counted on the leader's full barrier; commits are multicast to the same barrier offset in both CTAs.
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_addr` (a shared::cta address) in the CTA of rank `rank`
__device__ __forceinline__ uint32_t mapa_shared(uint32_t local_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint32_t leader_bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(leader_bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_tf32_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive (once all MMAs issued so far have completed) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t) 3) : "memory");
}

// ------------------------------------------------------------------------------------ descriptors
// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start>>4 [0,14), LBO>>4 [16,30),
// SBO>>4 [32,46), version=1 [46,48), layout type [61,64)
constexpr uint64_t LAYOUT_SW128 = 2;          // K-major operands: 128B swizzle, 16B atoms
constexpr uint64_t LAYOUT_SW128_BASE32B = 1;  // MN-major tf32 operands: 128B swizzle, 32B atoms
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint64_t layout) {
    return (uint64_t) ((saddr & 0x3FFFF) >> 4) | ((uint64_t) ((lbo_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t) ((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46) | (layout << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor) for kind::tf32, fp32 accumulate
__host__ __device__ constexpr uint32_t make_idesc(int m, int n, int a_mn_major, int b_mn_major) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t) a_mn_major << 15) | ((uint32_t) b_mn_major << 16) |
           ((uint32_t) (n >> 3) << 17) | ((uint32_t) (m >> 4) << 24);
}

struct EpiParams {
    const float* bias;      // [N] or null
    int relu;
    float keep;             // dropout keep prob, 1 = off
    uint64_t seed;
    const float* mask;      // [M, ldmask] or null
    int64_t ldmask;
    float mask_scale;
    int round_out;          // round outputs to tf32 (they feed the next tensor-core GEMM)
    int col0;               // first global output column of this CTA's slab (dropout hash index)
    int n_total;            // row pitch of the dropout hash index (the full N)
    int dbg;                // experiments only (CTR_GEMM_DBG): 1 skip global stores, 2 skip MMAs, 4 skip TMA loads, 16 direct 256-bit epilogue,
                            // 32 fast epilogue without its stores, 64 no epilogue at all (persistent kernel)
    unsigned long long* trace;   // CTR_GEMM_DBG & 8: per-CTA phase timestamps (8 slots per CTA)
    unsigned* sm_slots;     // split2 stagger: per-SM arrival counters (zeroed before the launch), or null
    unsigned stagger_ns;    // the second CTA to arrive on an SM starts this much later (first wave only)
    int tma_lanes;          // lanes of the producer warp that issue TMA loads (see tma_lanes() on the host side)
    // ReLU sign bits (ctr_mlp_fwd_bits / ctr_mlp_bwd_data_bits): one 32-bit word per (lane, 32 x 32 sub-tile) in the
    // epilogue's own thread <-> element mapping: nibble j of lane l = columns 4*(l&7)..+3 of row (l>>3) + 4*j of the
    // sub-tile.  Word index = ((row / 32) * bits_chunks + bits_c0 + c) * 32 + lane, c = the sub-tile's index inside the
    // CTA's unit / slab.  The producing and the consuming GEMM have the same (M, N), hence the same kernel and geometry.
    uint32_t* bits_out;     // forward: written when non-null (1 = stored activation > 0)
    const uint32_t* bits_in;   // data gradient: the mask, instead of `mask`
    int bits_chunks;        // sub-tile slots per 32-row block (host: launch_tn)
    int bits_c0;            // first slot of this CTA's unit / slab (set by the kernel)
};

__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// block 0 only: fine-grained slots after the per-CTA area (producer: 0.., MMA issuer: 64.., epilogue warp 0: 128..)
#define TRACE_FINE(slot) do { if (ep.trace && blockIdx.x == 0 && (slot) < 512) ep.trace[4096 * 8 + (slot)] = gtimer(); } while (0)
#define TRACE(slot) do { if (ep.trace && blockIdx.x < 4096) ep.trace[blockIdx.x * 8 + (slot)] = gtimer(); } while (0)

__device__ __forceinline__ float dropout_u01(uint64_t seed, uint64_t idx) {   // same hash as the FP32 path
    uint64_t z = seed + idx * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (float) (z >> 40) * (1.0f / 16777216.0f);
}

struct alignas(8) PipeBars {
    uint64_t full[6];
    uint64_t empty[6];
    uint64_t tmem_full;
    uint32_t tmem_base;
    uint32_t pad;
};

}  // namespace

namespace {

// ---- epilogue shared by both kernels: TMEM -> registers -> (bias / ReLU / dropout / mask) -> global.
// 8 warps: warp e owns TMEM lane quarter (e & 3) and every second 32-column chunk starting at (e >> 2).
// All per-element work stays in registers (fully unrolled, compile-time flags); the mask / bias loads
// of a chunk are issued before the TMEM load is waited on.
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// One 32x32 sub-tile per step: TMEM -> registers (lane = row) -> this warp's padded smem tile -> registers
// (8 lanes = one 128-byte row segment) -> bias / ReLU / mask / round / dropout -> global.  The transposition makes
// every global access of the epilogue (the stores AND the mask loads) a fully coalesced 128-bit access: a
// warp-wide store covers 4 complete 128-byte lines instead of 16 bytes in each of 32 different lines.
// `stage` is pipeline shared memory, free once tmem_full has fired (every TMA write has landed and every MMA
// that read it has completed).  Row pitch 36 floats keeps both the lane=row 128-bit writes and the 8-lanes-per-row
// 128-bit reads bank-conflict free.
constexpr int EPI_PITCH = 36;
constexpr int EPI_TILE_FLOATS = 32 * EPI_PITCH;

// explicit shared-space accesses: through a generic pointer the compiler emits LD/ST (generic), cannot tell the tile
// from the global output, and serialises every tile read behind the previous global store
__device__ __forceinline__ void sts_f4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}

// round-to-nearest (ties away) fp32 -> tf32 for finite values: what cvt.rna.tf32.f32 returns, in two integer ops
// (ptxas expands the cvt into a longer sequence with inf/nan handling)
__device__ __forceinline__ float round_tf32_fast(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }

// fp32 -> tf32 (round to nearest even) in ONE instruction (F2FP.TF32), optionally with the ReLU folded in
__device__ __forceinline__ float cvt_tf32(float x) {
    uint32_t r;
    asm("cvt.rn.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
__device__ __forceinline__ float cvt_tf32_relu(float x) {
    uint32_t r;
    asm("cvt.rn.relu.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

__device__ __forceinline__ int rbase_of(int m0, int ewarp) { return m0 + (ewarp & 3) * 32; }

template <bool BIAS, bool RELU, int MASK, int EW>
__device__ __forceinline__ void epilogue_rows(uint32_t tmem_base, int ewarp, int lane, int m0, int M, int N, float* C,
                                              int64_t ldc, const EpiParams& ep, float* stage, uint32_t sbias) {
    const int q = ewarp & 3, half = ewarp >> 2;     // TMEM lane quarter; chunk phase (stride EW / 4)
    const uint32_t tile = smem_u32(stage) + (uint32_t) (ewarp * EPI_TILE_FLOATS * 4);
    const bool vec_ok = (ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(C) & 15) == 0);
    const bool mvec_ok = MASK == 1 && (ep.ldmask % 4 == 0) && ((reinterpret_cast<uintptr_t>(ep.mask) & 15) == 0);
    const bool dropout = ep.keep < 1.f;
    const size_t bits_row = (size_t) (rbase_of(m0, ewarp) >> 5) * (size_t) ep.bits_chunks + (size_t) ep.bits_c0;
    const uint32_t lane_addr = tmem_base + ((uint32_t) (q * 32) << 16);
    const int nchunks = (N + 31) / 32;
    const int rsub = lane >> 3;            // row inside a group of 4 rows
    const int cc = (lane & 7) * 4;         // first of this lane's 4 columns inside the chunk
    const int rbase = m0 + q * 32;
    for (int c = half; c < nchunks; c += EW / 4) {
        const int c0 = c * 32;
        uint32_t r[32];
        tmem_ld32_issue(lane_addr + (uint32_t) c0, r);   // .sync.aligned: every lane takes part, valid row or not
        const int col = c0 + cc;
        const bool col4 = col + 4 <= N;
        float4 bs = make_float4(0.f, 0.f, 0.f, 0.f);
        if (BIAS) {     // staged in shared memory by the epilogue warps while the mainloop ran (zero-padded to 512)
            bs = lds_f4(sbias + (uint32_t) (col * 4));
        }
        float4 mk[8];
        uint32_t mbits = 0, obits = 0;
        if (MASK == 2) mbits = __ldg(ep.bits_in + (bits_row + (size_t) c) * 32 + lane);
        if (MASK == 1) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int row = rbase + rsub + 4 * j;
                const float* mp = ep.mask + (int64_t) row * ep.ldmask + col;
                if (row < M && col4 && mvec_ok) mk[j] = ctr_ldg_f4_stream(mp);
                else {
                    mk[j].x = (row < M && col + 0 < N) ? mp[0] : 0.f;
                    mk[j].y = (row < M && col + 1 < N) ? mp[1] : 0.f;
                    mk[j].z = (row < M && col + 2 < N) ? mp[2] : 0.f;
                    mk[j].w = (row < M && col + 3 < N) ? mp[3] : 0.f;
                }
            }
        }
        tmem_ld_wait();
        if (ewarp == 0 && lane == 0) TRACE_FINE(128 + 4 * c);
#pragma unroll
        for (int i = 0; i < 8; ++i)
            sts_f4(tile + (uint32_t) ((lane * EPI_PITCH + 4 * i) * 4), __uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]),
                   __uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]));
        __syncwarp();
        float4 tv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) tv[j] = lds_f4(tile + (uint32_t) (((rsub + 4 * j) * EPI_PITCH + cc) * 4));
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int rr = rsub + 4 * j;
            const int row = rbase + rr;
            float4 v = tv[j];
            if (BIAS) { v.x += bs.x; v.y += bs.y; v.z += bs.z; v.w += bs.w; }
            if (RELU) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
            if (MASK == 1) {
                v.x = mk[j].x > 0.f ? v.x * ep.mask_scale : 0.f;
                v.y = mk[j].y > 0.f ? v.y * ep.mask_scale : 0.f;
                v.z = mk[j].z > 0.f ? v.z * ep.mask_scale : 0.f;
                v.w = mk[j].w > 0.f ? v.w * ep.mask_scale : 0.f;
            }
            if (MASK == 2) {
                const uint32_t nib = mbits >> (4 * j);
                v.x = (nib & 1u) ? v.x * ep.mask_scale : 0.f;
                v.y = (nib & 2u) ? v.y * ep.mask_scale : 0.f;
                v.z = (nib & 4u) ? v.z * ep.mask_scale : 0.f;
                v.w = (nib & 8u) ? v.w * ep.mask_scale : 0.f;
            }
            if (ep.round_out) { v.x = cvt_tf32(v.x); v.y = cvt_tf32(v.y); v.z = cvt_tf32(v.z); v.w = cvt_tf32(v.w); }
            if (dropout) {
                const uint64_t e0 = (uint64_t) row * (uint64_t) (ep.n_total > 0 ? ep.n_total : N) + (uint64_t) (ep.col0 + col);
                v.x = dropout_u01(ep.seed, e0 + 0) < ep.keep ? v.x / ep.keep : 0.f;
                v.y = dropout_u01(ep.seed, e0 + 1) < ep.keep ? v.y / ep.keep : 0.f;
                v.z = dropout_u01(ep.seed, e0 + 2) < ep.keep ? v.z / ep.keep : 0.f;
                v.w = dropout_u01(ep.seed, e0 + 3) < ep.keep ? v.w / ep.keep : 0.f;
            }
            if (ep.bits_out)       // what the data gradient masks with: the STORED activation (after dropout) > 0
                obits |= ((col + 0 < N && v.x > 0.f ? 1u : 0u) | (col + 1 < N && v.y > 0.f ? 2u : 0u) |
                          (col + 2 < N && v.z > 0.f ? 4u : 0u) | (col + 3 < N && v.w > 0.f ? 8u : 0u)) << (4 * j);
            if (row < M && !(ep.dbg & 1)) {
                float* dst = C + (int64_t) row * ldc + col;
                if (vec_ok && col4) *reinterpret_cast<float4*>(dst) = v;
                else {
                    if (col + 0 < N) dst[0] = v.x;
                    if (col + 1 < N) dst[1] = v.y;
                    if (col + 2 < N) dst[2] = v.z;
                    if (col + 3 < N) dst[3] = v.w;
                }
            }
        }
        if (ep.bits_out) ep.bits_out[(bits_row + (size_t) c) * 32 + lane] = obits;
        if (ewarp == 0 && lane == 0) TRACE_FINE(128 + 4 * c + 1);
        __syncwarp();      // the tile is rewritten by the next chunk
    }
}

// The same epilogue for the common case -- every row of the tile valid, 16-byte aligned vector accesses, N % 4 == 0,
// rounded outputs, no dropout: ~115 instructions per 32x32 sub-tile instead of ~330.  The epilogue is issue-bound (two
// warps per scheduler, each a dependent chain), so its duration is its instruction count: bounds checks, per-row
// 64-bit address arithmetic and the 3-instruction ReLU + round were two thirds of it.  The ReLU mask of the data-gradient
// GEMM is fetched one sub-tile ahead (its DRAM latency was the whole epilogue: 16 us per tile against 5.7 without it).
// Tried and dropped: a transposition-free epilogue on tcgen05.ld.16x256b fragments (32-byte runs per thread quad, 64-bit
// stores / mask loads) -- it frees the shared-memory port but the 8-sectors-per-instruction accesses made the
// epilogue 1.5x (stores) to 2x (mask loads) slower.
template <bool BIAS, bool RELU, int MASK, bool BITS_OUT, int EW, bool ROUND = true, bool MTAIL = false>
__device__ __forceinline__ void epilogue_rows_fast(uint32_t tmem_base, int ewarp, int lane, int m0, int M, int N, float* C,
                                                   int64_t ldc, const EpiParams& ep, float* stage, uint32_t sbias) {
    const int q = ewarp & 3, half = ewarp >> 2;
    const uint32_t tile = smem_u32(stage) + (uint32_t) (ewarp * EPI_TILE_FLOATS * 4);
    const uint32_t lane_addr = tmem_base + ((uint32_t) (q * 32) << 16);
    const int nchunks = (N + 31) / 32;
    const int rsub = lane >> 3, cc = (lane & 7) * 4;
    const int row0 = m0 + q * 32 + rsub;
    // MTAIL: rows row0 + 4 j, j < jmax, exist (the last tile of the matrix may be partial); else the tile is whole
    const int jmax = !MTAIL ? 8 : (M - row0 >= 32 ? 8 : (M > row0 ? (M - row0 + 3) >> 2 : 0));
    const uint32_t st_addr = tile + (uint32_t) (lane * EPI_PITCH * 4);
    const uint32_t ld_addr = tile + (uint32_t) ((rsub * EPI_PITCH + cc) * 4);
    char* dst_row = reinterpret_cast<char*>(C + (int64_t) row0 * ldc + cc);
    const char* msk_row = MASK == 1 ? reinterpret_cast<const char*>(ep.mask + (int64_t) row0 * ep.ldmask + cc) : nullptr;
    // sign-bit words of this warp's 32-row block (see EpiParams): one coalesced 128-byte access per sub-tile
    const size_t bits_row = ((size_t) ((m0 >> 5) + q) * (size_t) ep.bits_chunks + (size_t) ep.bits_c0) * 32 + (size_t) lane;
    uint32_t mbits = 0, mbits_next = 0;
    const int64_t dstep = 4 * ldc * 4, mstep = 4 * ep.ldmask * 4;         // 4 rows down, in bytes
    const float scale = ep.mask_scale;
    float4 mk[8], mkn[8];
    auto load_mask = [&](int c, float4* m) {
        const bool ok = c < nchunks && c * 32 + cc < N;
        const char* mp = msk_row + (int64_t) c * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            m[j] = (ok && j < jmax) ? ctr_ldg_f4_stream(reinterpret_cast<const float*>(mp)) : make_float4(0.f, 0.f, 0.f, 0.f);
            mp += mstep;
        }
    };
    if (MASK == 1) load_mask(half, mkn);
    if (MASK == 2 && half < nchunks) mbits_next = __ldg(ep.bits_in + bits_row + (size_t) half * 32);
    for (int c = half; c < nchunks; c += EW / 4) {
        const int c0 = c * 32;
        uint32_t r[32];
        tmem_ld32_issue(lane_addr + (uint32_t) c0, r);
        const bool col_ok = c0 + cc < N;                  // N % 4 == 0: a lane's four columns are valid together
        float4 bs = make_float4(0.f, 0.f, 0.f, 0.f);
        if (BIAS) bs = lds_f4(sbias + (uint32_t) ((c0 + cc) * 4));
        if (MASK == 1) {
#pragma unroll
            for (int j = 0; j < 8; ++j) mk[j] = mkn[j];
            load_mask(c + EW / 4, mkn);                   // the next sub-tile's mask, in flight under this one
        }
        if (MASK == 2) {
            mbits = mbits_next;
            if (c + EW / 4 < nchunks) mbits_next = __ldg(ep.bits_in + bits_row + (size_t) (c + EW / 4) * 32);
        }
        uint32_t obits = 0;
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i)
            sts_f4(st_addr + (uint32_t) (16 * i), __uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]),
                   __uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]));
        __syncwarp();
        float4 tv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) tv[j] = lds_f4(ld_addr + (uint32_t) (4 * j * EPI_PITCH * 4));
        char* dp = dst_row + (int64_t) c0 * 4;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float4 v = tv[j];
            if (BIAS) { v.x += bs.x; v.y += bs.y; v.z += bs.z; v.w += bs.w; }
            if (MASK == 1) {
                v.x = mk[j].x > 0.f ? v.x * scale : 0.f; v.y = mk[j].y > 0.f ? v.y * scale : 0.f;
                v.z = mk[j].z > 0.f ? v.z * scale : 0.f; v.w = mk[j].w > 0.f ? v.w * scale : 0.f;
            }
            if (MASK == 2) {
                v.x = (mbits & (1u << (4 * j))) ? v.x * scale : 0.f; v.y = (mbits & (2u << (4 * j))) ? v.y * scale : 0.f;
                v.z = (mbits & (4u << (4 * j))) ? v.z * scale : 0.f; v.w = (mbits & (8u << (4 * j))) ? v.w * scale : 0.f;
            }
            if (RELU) { v.x = cvt_tf32_relu(v.x); v.y = cvt_tf32_relu(v.y); v.z = cvt_tf32_relu(v.z); v.w = cvt_tf32_relu(v.w); }
            else if (ROUND) { v.x = cvt_tf32(v.x); v.y = cvt_tf32(v.y); v.z = cvt_tf32(v.z); v.w = cvt_tf32(v.w); }
            if (BITS_OUT && col_ok) {
                // the stored value is >= +0 after the ReLU, so "> 0" is "bit pattern != 0": min(u, 1) is the bit
                obits |= (min(__float_as_uint(v.x), 1u) | (min(__float_as_uint(v.y), 1u) << 1) |
                          (min(__float_as_uint(v.z), 1u) << 2) | (min(__float_as_uint(v.w), 1u) << 3)) << (4 * j);
            }
            if (col_ok && j < jmax && !(ep.dbg & 32)) {
                if (ep.dbg & 128) __stcs(reinterpret_cast<float4*>(dp), v);       // experiment: streaming (evict-first) stores
                else *reinterpret_cast<float4*>(dp) = v;
            }
            dp += dstep;
        }
        if (BITS_OUT) ep.bits_out[bits_row + (size_t) c * 32] = obits;
        __syncwarp();      // the tile is rewritten by the next chunk
    }
}

// EXPERIMENT (CTR_GEMM_DBG & 16, default off, not yet run on a GPU): the same epilogue without the shared-memory
// transposition.  Blackwell has 256-bit global accesses (STG.E.ENL2.256 / LDG.E.ENL2.256): with the accumulator in the
// 32x32b layout (lane = row, 32 consecutive columns per lane) a lane stores one complete 32-byte sector per instruction,
// so every sector is written whole although a warp touches 32 different rows; four instructions complete each 128-byte
// line.  Frees the shared-memory port (the transposition was ~28 B/clk of the 128 the MMAs and TMA fills compete for) at
// the price of 4x the L2 requests per byte of the coalesced stores.  The 16x256b attempt above had the same request
// count but 64-bit accesses, i.e. 4x the instructions as well.
__device__ __forceinline__ void stg_f8(float* p, const float* v) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
                 "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}
__device__ __forceinline__ void ldg_f8_stream(const float* p, float* v) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p));
}

template <bool BIAS, bool RELU, bool MASK, int EW>
__device__ __forceinline__ void epilogue_rows_direct(uint32_t tmem_base, int ewarp, int lane, int m0, int N, float* C,
                                                     int64_t ldc, const EpiParams& ep, uint32_t sbias) {
    const int q = ewarp & 3, half = ewarp >> 2;
    const uint32_t lane_addr = tmem_base + ((uint32_t) (q * 32) << 16);
    const int nchunks = (N + 31) / 32;
    const int64_t row = m0 + q * 32 + lane;
    float* dst_row = C + row * ldc;
    const float* msk_row = MASK ? ep.mask + row * ep.ldmask : nullptr;
    const float scale = ep.mask_scale;
    float mk[32];                                         // ONE buffer: group g of the next sub-tile is requested as soon as
                                                          // group g of this one has been consumed (a second buffer spills
                                                          // under the persistent kernel's 128-register cap)
    auto load_group = [&](int c, int g) {
        if (c < nchunks && c * 32 + 8 * g < N) ldg_f8_stream(msk_row + c * 32 + 8 * g, mk + 8 * g);
    };
    if (MASK) {
#pragma unroll
        for (int g = 0; g < 4; ++g) load_group(half, g);
    }
    for (int c = half; c < nchunks; c += EW / 4) {
        const int c0 = c * 32;
        uint32_t r[32];
        tmem_ld32_issue(lane_addr + (uint32_t) c0, r);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if (c0 + 8 * g < N) {                         // N % 8 == 0: a group of eight columns is valid together
                float v[8];
                float4 b0 = make_float4(0.f, 0.f, 0.f, 0.f), b1 = b0;
                if (BIAS) {                               // every lane reads the same address: a broadcast, no bank conflict
                    b0 = lds_f4(sbias + (uint32_t) ((c0 + 8 * g) * 4));
                    b1 = lds_f4(sbias + (uint32_t) ((c0 + 8 * g + 4) * 4));
                }
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float x = __uint_as_float(r[8 * g + i]);
                    if (BIAS) x += bb[i];
                    if (MASK) x = mk[8 * g + i] > 0.f ? x * scale : 0.f;
                    v[i] = RELU ? cvt_tf32_relu(x) : cvt_tf32(x);
                }
                stg_f8(dst_row + c0 + 8 * g, v);
            }
            if (MASK) load_group(c + EW / 4, g);
        }
    }
}

// shared memory: [ pipeline stages, reused as epilogue staging | PipeBars | staged bias ]
__host__ __device__ constexpr size_t pipe_region(size_t stage_total, int ew) {
    return stage_total > (size_t) ew * EPI_TILE_FLOATS * 4 ? stage_total : (size_t) ew * EPI_TILE_FLOATS * 4;
}
constexpr size_t smem_total(size_t stage_total, int ew) { return pipe_region(stage_total, ew) + sizeof(PipeBars) + SBIAS_BYTES + 1024; }

// what the epilogue warps do BEFORE the accumulator is ready (they are idle during the mainloop): stage the bias of
// this CTA's columns in shared memory (an L2 prefetch of the dgrad's mask tile was tried here: it slowed the kernel down)
template <int EW>
__device__ __forceinline__ void epilogue_prepare(int m0, int M, int N, const EpiParams& ep, float* sbias) {
    const int tid = threadIdx.x;                   // epilogue warps are warps [0, EW)
    for (int i = tid; i < SBIAS_BYTES / 4; i += EW * 32) sbias[i] = (ep.bias && i < N) ? __ldg(ep.bias + i) : 0.f;
    asm volatile("bar.sync 1, %0;" ::"r"(EW * 32) : "memory");      // epilogue warps only
}

template <int EW, bool DIRECT = false>
__device__ __forceinline__ void epilogue_dispatch(uint32_t tmem_base, int ewarp, int lane, int m0, int M, int N, float* C,
                                                  int64_t ldc, const EpiParams& ep, float* stage, uint32_t sbias) {
    const bool fast = m0 + BLOCK_M <= M && (N & 3) == 0 && (ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0 &&
                      ep.round_out && !(ep.keep < 1.f) && !(ep.dbg & 1) &&
                      (!ep.mask || ((ep.ldmask & 3) == 0 && (reinterpret_cast<uintptr_t>(ep.mask) & 15) == 0));
    const bool bits = ep.bits_in != nullptr || ep.bits_out != nullptr;
    if (DIRECT && fast && !bits && (N & 7) == 0 && (ldc & 7) == 0 && (reinterpret_cast<uintptr_t>(C) & 31) == 0 &&
        (!ep.mask || ((ep.ldmask & 7) == 0 && (reinterpret_cast<uintptr_t>(ep.mask) & 31) == 0))) {
        if (ep.mask) epilogue_rows_direct<false, false, true, EW>(tmem_base, ewarp, lane, m0, N, C, ldc, ep, sbias);
        else if (ep.bias && ep.relu) epilogue_rows_direct<true, true, false, EW>(tmem_base, ewarp, lane, m0, N, C, ldc, ep, sbias);
        else if (ep.bias) epilogue_rows_direct<true, false, false, EW>(tmem_base, ewarp, lane, m0, N, C, ldc, ep, sbias);
        else if (ep.relu) epilogue_rows_direct<false, true, false, EW>(tmem_base, ewarp, lane, m0, N, C, ldc, ep, sbias);
        else epilogue_rows_direct<false, false, false, EW>(tmem_base, ewarp, lane, m0, N, C, ldc, ep, sbias);
        return;
    }
    // sign bits are written only behind a ReLU (the fast path's "bit pattern != 0" test relies on stored values >= +0)
    if (fast && !(ep.bits_out && !(ep.bias && ep.relu))) {
        if (ep.bits_in) epilogue_rows_fast<false, false, 2, false, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
        else if (ep.mask) epilogue_rows_fast<false, false, 1, false, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
        else if (ep.bias && ep.relu && ep.bits_out)
            epilogue_rows_fast<true, true, 0, true, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
        else if (ep.bias && ep.relu) epilogue_rows_fast<true, true, 0, false, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
        else if (ep.bias) epilogue_rows_fast<true, false, 0, false, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
        else if (ep.relu) epilogue_rows_fast<false, true, 0, false, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
        else epilogue_rows_fast<false, false, 0, false, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
        return;
    }
    if (ep.bits_in) epilogue_rows<false, false, 2, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
    else if (ep.mask) epilogue_rows<false, false, 1, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
    else if (ep.bias && ep.relu) epilogue_rows<true, true, 0, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
    else if (ep.bias) epilogue_rows<true, false, 0, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
    else if (ep.relu) epilogue_rows<false, true, 0, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
    else epilogue_rows<false, false, 0, EW>(tmem_base, ewarp, lane, m0, M, N, C, ldc, ep, stage, sbias);
}

__device__ __forceinline__ PipeBars* setup_cta(uint8_t* smem, size_t pipe_bytes, int nstages, uint32_t tmem_cols, int warp, int tma_warp) {
    PipeBars* bars = reinterpret_cast<PipeBars*>(smem + pipe_bytes);
    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        mbar_init(&bars->tmem_full, 1);
        fence_barrier_init();
    }
    if (warp == tma_warp) tmem_alloc(&bars->tmem_base, tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    return bars;
}

// ---- C[M,N] = A[M,R] * B[N,R]^T ; A, B reduction-major, 128B-swizzled K-major tiles -------------------
// MINB = 2: two CTAs share an SM (each <= 256 TMEM columns, <= ~113 KB shared memory), so one CTA's epilogue and
// pipeline fill run under the other's MMAs; blockIdx.y selects the CTA's slab of n_each * nmma output columns.
template <int MINB, int EW>
__global__ void __launch_bounds__(nthreads(EW), MINB)
gemm_tn_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N_total, int R,
               int n_each, int nmma, int nslab, int nstages, uint32_t tmem_cols, float* __restrict__ C, int64_t ldc,
               EpiParams ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int NB = n_each * nmma;
    // slab index fastest: the CTAs that share a 128-row A panel launch back to back, so the second read hits L2
    const int n0 = ((int) blockIdx.x % nslab) * NB;                   // first output column of this CTA
    const int N = (N_total - n0) < NB ? (N_total - n0) : NB;          // columns this CTA writes
    C += n0;
    if (ep.bias) ep.bias += n0;
    if (ep.mask) ep.mask += n0;
    ep.col0 = n0;
    ep.bits_c0 = ((int) blockIdx.x % nslab) * ((NB + 31) / 32);
    const int a_bytes = BLOCK_M * BLOCK_R * 4;
    const int stage_bytes = a_bytes + NB * BLOCK_R * 4;
    const int m0 = ((int) blockIdx.x / nslab) * BLOCK_M;
    const int num_kb = (R + BLOCK_R - 1) / BLOCK_R;
    if (threadIdx.x == 0) TRACE(0);
    if (MINB == 2 && ep.sm_slots) {
        // de-phase the two CTAs that share an SM: left alone they run in lockstep (both in the TMA-bound main loop, then
        // both in the store-bound epilogue); started half a tile apart, one's epilogue runs under the other's main loop,
        // and since a CTA's successor starts when it exits the offset persists for the rest of the grid
        if (threadIdx.x == 0) {
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            if (atomicAdd(ep.sm_slots + smid, 1u) == 1u) {
                const unsigned long long t0 = gtimer();
                while (gtimer() - t0 < ep.stagger_ns) __nanosleep(256);
            }
        }
        __syncthreads();
    }
    constexpr int TMA_WARP = EW, MMA_WARP = EW + 1;
    const size_t pipe_bytes = pipe_region((size_t) nstages * stage_bytes, EW);
    PipeBars* bars = setup_cta(smem, pipe_bytes, nstages, tmem_cols, warp, TMA_WARP);
    float* sbias = reinterpret_cast<float*>(smem + pipe_bytes + sizeof(PipeBars));
    const uint32_t tmem_base = bars->tmem_base;
    if (threadIdx.x == 0) TRACE(1);

    if (warp == TMA_WARP) {
        if (elect_one()) {
            prefetch_tmap(&tmA);
            prefetch_tmap(&tmB);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % nstages;
                const uint32_t ph = (uint32_t) (kb / nstages) & 1u;
                mbar_wait(&bars->empty[s], ph ^ 1u);
                TRACE_FINE(kb);
                uint8_t* sa = smem + (size_t) s * stage_bytes;
                if (ep.dbg & 4) { mbar_expect_tx(&bars->full[s], 0u); continue; }
                mbar_expect_tx(&bars->full[s], (uint32_t) stage_bytes);
                tma_load_2d(sa, &tmA, &bars->full[s], kb * BLOCK_R, m0);
                for (int i = 0; i < nmma; ++i)
                    tma_load_2d(sa + a_bytes + (size_t) i * n_each * BLOCK_R * 4, &tmB, &bars->full[s], kb * BLOCK_R, n0 + i * n_each);
            }
        }
    } else if (warp == MMA_WARP) {
        if (elect_one()) {
            const uint32_t idesc = make_idesc(BLOCK_M, n_each, 0, 0);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % nstages;
                const uint32_t ph = (uint32_t) (kb / nstages) & 1u;
                mbar_wait(&bars->full[s], ph);
                tc_fence_after();
                TRACE_FINE(64 + kb);
                const uint32_t a_base = smem_u32(smem + (size_t) s * stage_bytes);
                const uint32_t b_base = a_base + a_bytes;
#pragma unroll
                for (int k = 0; k < BLOCK_R / UMMA_K; ++k) {
                    const uint64_t adesc = make_desc(a_base + k * UMMA_K * 4, 16, 8 * BLOCK_R * 4, LAYOUT_SW128);
                    for (int i = 0; i < nmma; ++i) {
                        const uint64_t bdesc = make_desc(b_base + i * n_each * BLOCK_R * 4 + k * UMMA_K * 4, 16, 8 * BLOCK_R * 4, LAYOUT_SW128);
                        if (!(ep.dbg & 2)) umma_tf32(tmem_base + (uint32_t) (i * n_each), adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                }
                umma_commit(&bars->empty[s]);     // frees the smem stage once these MMAs have read it
            }
            umma_commit(&bars->tmem_full);        // accumulator complete
            TRACE(2);
        }
    } else {
        epilogue_prepare<EW>(m0, M, N, ep, sbias);
        mbar_wait(&bars->tmem_full, 0);
        tc_fence_after();
        if (threadIdx.x == 0) TRACE(3);
        epilogue_dispatch<EW>(tmem_base, warp, lane, m0, M, N, C, ldc, ep, reinterpret_cast<float*>(smem), smem_u32(sbias));
        if (threadIdx.x == 0) TRACE(4);
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) TRACE(5);
    if (warp == TMA_WARP) tmem_dealloc(tmem_base, tmem_cols);
    if (threadIdx.x == TMA_WARP * 32) TRACE(6);
}

// ---- C[M,N] = A[M,R] * B[N,R]^T on CTA pairs (see the cta_group::2 notes above).  CTA rank r of pair p owns output
// rows [(2p + r) * 128, +128) and stages rows [i*n_each + r*n_each/2, +n_each/2) of B for MMA i.
template <int EW>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(nthreads(EW), 1)
gemm_tn_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N, int R,
                    int n_each, int nmma, int nstages, uint32_t tmem_cols, float* __restrict__ C, int64_t ldc, EpiParams ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int n_half = n_each / 2;
    const int a_bytes = BLOCK_M * BLOCK_R * 4;
    const int b_bytes = n_half * BLOCK_R * 4;                 // one MMA's B half
    const int stage_bytes = a_bytes + nmma * b_bytes;
    const int m0 = blockIdx.x * BLOCK_M;
    const int num_kb = (R + BLOCK_R - 1) / BLOCK_R;
    constexpr int TMA_WARP = EW, MMA_WARP = EW + 1;
    const size_t pipe_bytes = pipe_region((size_t) nstages * stage_bytes, EW);
    PipeBars* bars = reinterpret_cast<PipeBars*>(smem + pipe_bytes);
    float* sbias = reinterpret_cast<float*>(smem + pipe_bytes + sizeof(PipeBars));
    if (threadIdx.x == 0) TRACE(0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        mbar_init(&bars->tmem_full, 1);
        fence_barrier_init();
    }
    cluster_sync_all();                                      // both CTAs' barriers exist before anyone signals them
    if (warp == TMA_WARP) tmem_alloc_pair(&bars->tmem_base, tmem_cols);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    if (threadIdx.x == 0) TRACE(1);

    if (warp == TMA_WARP) {
        if (elect_one()) {
            prefetch_tmap(&tmA);
            prefetch_tmap(&tmB);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % nstages;
                const uint32_t ph = (uint32_t) (kb / nstages) & 1u;
                mbar_wait(&bars->empty[s], ph ^ 1u);           // my copy of the stage is free (multicast commit)
                TRACE_FINE(kb);
                uint8_t* sa = smem + (size_t) s * stage_bytes;
                if (rank == 0) mbar_expect_tx(&bars->full[s], (uint32_t) (2 * stage_bytes));   // both CTAs' bytes
                const uint32_t lbar = mapa_shared(smem_u32(&bars->full[s]), 0);
                tma_load_2d_pair(sa, &tmA, lbar, kb * BLOCK_R, m0);
                for (int i = 0; i < nmma; ++i)
                    tma_load_2d_pair(sa + a_bytes + (size_t) i * b_bytes, &tmB, lbar, kb * BLOCK_R,
                                     i * n_each + (int) rank * n_half);
            }
        }
    } else if (warp == MMA_WARP) {
        if (rank == 0 && elect_one()) {
            const uint32_t idesc = make_idesc(2 * BLOCK_M, n_each, 0, 0);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % nstages;
                const uint32_t ph = (uint32_t) (kb / nstages) & 1u;
                mbar_wait(&bars->full[s], ph);
                tc_fence_after();
                TRACE_FINE(64 + kb);
                const uint32_t a_base = smem_u32(smem + (size_t) s * stage_bytes);
                const uint32_t b_base = a_base + a_bytes;
#pragma unroll
                for (int k = 0; k < BLOCK_R / UMMA_K; ++k) {
                    const uint64_t adesc = make_desc(a_base + k * UMMA_K * 4, 16, 8 * BLOCK_R * 4, LAYOUT_SW128);
                    for (int i = 0; i < nmma; ++i) {
                        const uint64_t bdesc = make_desc(b_base + i * b_bytes + k * UMMA_K * 4, 16, 8 * BLOCK_R * 4, LAYOUT_SW128);
                        umma_tf32_pair(tmem_base + (uint32_t) (i * n_each), adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                }
                umma_commit_pair(&bars->empty[s]);     // frees the stage in BOTH CTAs
            }
            umma_commit_pair(&bars->tmem_full);        // accumulators complete in both CTAs
            TRACE(2);
        }
    } else {
        epilogue_prepare<EW>(m0, M, N, ep, sbias);
        mbar_wait(&bars->tmem_full, 0);
        tc_fence_after();
        if (threadIdx.x == 0) TRACE(3);
        epilogue_dispatch<EW>(tmem_base, warp, lane, m0, M, N, C, ldc, ep, reinterpret_cast<float*>(smem), smem_u32(sbias));
        if (threadIdx.x == 0) TRACE(4);
    }
    tc_fence_before();
    cluster_sync_all();                                   // neither CTA frees TMEM / exits while its peer still runs
    if (threadIdx.x == 0) TRACE(5);
    if (warp == TMA_WARP) tmem_dealloc_pair(tmem_base, tmem_cols);
    if (threadIdx.x == TMA_WARP * 32) TRACE(6);
}

// ---- persistent CTA pairs with a double-buffered accumulator.  A cluster walks pair tiles (256 rows) cid, cid + n_clusters,
// ...; each tile is nmma work units of n_each (<= 256) output columns, and unit u accumulates in TMEM columns
// [(u & 1) * 256, +n_each), so the epilogue of unit u (8 warps per CTA: TMEM -> registers -> bias / ReLU / mask / round
// -> coalesced stores) runs under the MMAs of unit u + 1.  Barriers, TMEM and the staged bias live for the whole
// launch; the stage ring runs on across units, so the TMA warps are always a few k-blocks ahead of the MMA issuer.
// Per k-block a CTA stages its 128 rows of A (16 KB) and n_each / 2 rows of B (the peer CTA holds the other half): A is
// streamed once per unit, the second time from L2.  Why this shape: the one-tile-per-CTA kernels above run main loop
// and epilogue back to back (tests/trace_gemm.py: 12 + 7.4 us per tile, in lockstep across the GPU), and an N = 400
// fp32 accumulator fills 400 of the 512 TMEM columns, so only half-N units leave room for a second buffer.
struct alignas(16) PairBars {      // sizeof % 16 == 0: the staged bias behind it is read with 128-bit loads
    uint64_t full[6];
    uint64_t empty[6];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint32_t tmem_base;
    uint32_t pad[3];
};
static_assert(sizeof(PairBars) % 16 == 0, "PairBars must keep the bias 16-byte aligned");
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
constexpr size_t pair_persistent_smem(size_t stage_total, int ew) {
    return stage_total + (size_t) ew * EPI_TILE_FLOATS * 4 + sizeof(PairBars) + SBIAS_BYTES + 1024;
}

// 128 registers per thread (not the 204 a lone 320-thread CTA could have): the CTA owns its SM for the whole launch, and
// the side stream's id sort (radix scatter: 256 threads x 96 registers, 10 KB of shared memory) must still fit next to it,
// or that work only runs in the gaps between GEMMs and lands on the step's critical path
template <int EW, bool DIRECT = false>
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(128)
gemm_tn_pair_persistent_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N,
                               int R, int n_each, int nmma, int nstages, float* __restrict__ C, int64_t ldc, EpiParams ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int cid = (int) blockIdx.x >> 1, ncl = (int) gridDim.x >> 1;
    const int n_tiles = (M + 2 * BLOCK_M - 1) / (2 * BLOCK_M);
    const int my_tiles = cid < n_tiles ? (n_tiles - cid + ncl - 1) / ncl : 0;
    const uint32_t n_units = (uint32_t) (my_tiles * nmma);
    const int n_half = n_each / 2;
    const int a_bytes = BLOCK_M * BLOCK_R * 4;
    const int b_bytes = n_half * BLOCK_R * 4;
    const int stage_bytes = a_bytes + b_bytes;
    const int num_kb = (R + BLOCK_R - 1) / BLOCK_R;
    constexpr int TMA_WARP = EW, MMA_WARP = EW + 1;
    constexpr uint32_t TMEM_COLS = 512, BUF_COLS = 256;
    const size_t pipe_bytes = (size_t) nstages * stage_bytes;
    float* estage = reinterpret_cast<float*>(smem + pipe_bytes);
    PairBars* bars = reinterpret_cast<PairBars*>(smem + pipe_bytes + (size_t) EW * EPI_TILE_FLOATS * 4);
    float* sbias = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + sizeof(PairBars));
    if (threadIdx.x == 0) TRACE(0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&bars->tmem_full[b], 1);
            mbar_init(&bars->tmem_empty[b], 2 * EW);       // every epilogue warp of both CTAs
        }
        fence_barrier_init();
    }
    cluster_sync_all();                                      // both CTAs' barriers exist before anyone signals them
    if (warp == TMA_WARP) tmem_alloc_pair(&bars->tmem_base, TMEM_COLS);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    if (threadIdx.x == 0) TRACE(1);

    if (warp == TMA_WARP) {
        // One thread gets a TMA tile load through every ~0.3 us whatever its size (tools/probe_tma_fill.cu,
        // profiles/r02j_probe_tma_fill.txt: 43 GB/s per SM from one issuing thread, 124 from four, lanes of one warp or
        // separate warps alike), and the two loads of a k-block issued by one thread WERE this kernel's main loop
        // (0.64 us per k-block against ~0.2 us of MMAs).  So the k-blocks are dealt round-robin to `nl` lanes, each
        // running its own wait -> expect -> load loop; the stages still fill in ring order per lane, and the MMA issuer
        // consumes them in order whichever lane filled them.
        const int nl = ep.tma_lanes < 1 ? 1 : (ep.tma_lanes > nstages ? nstages : ep.tma_lanes);
        if (lane < nl) {
            if (lane == 0) {
                prefetch_tmap(&tmA);
                prefetch_tmap(&tmB);
            }
            const uint32_t total = n_units * (uint32_t) num_kb;
            for (uint32_t it = (uint32_t) lane; it < total; it += (uint32_t) nl) {
                const uint32_t u = it / (uint32_t) num_kb;
                const int kb = (int) (it - u * (uint32_t) num_kb);
                const int tile = cid + (int) (u / (uint32_t) nmma) * ncl, hh = (int) (u % (uint32_t) nmma);
                const int m0 = (tile * 2 + (int) rank) * BLOCK_M;
                const int s = (int) (it % (uint32_t) nstages);
                const uint32_t ph = (it / (uint32_t) nstages) & 1u;
                mbar_wait(&bars->empty[s], ph ^ 1u);       // my copy of the stage is free (multicast commit)
                uint8_t* sa = smem + (size_t) s * stage_bytes;
                if (ep.dbg & 4) {                          // timing experiment: no loads, the MMAs read whatever is there
                    if (rank == 0) mbar_expect_tx(&bars->full[s], 0u);
                    continue;
                }
                if (rank == 0) mbar_expect_tx(&bars->full[s], (uint32_t) (2 * stage_bytes));   // both CTAs' bytes
                const uint32_t lbar = mapa_shared(smem_u32(&bars->full[s]), 0);
                tma_load_2d_pair(sa, &tmA, lbar, kb * BLOCK_R, m0);
                tma_load_2d_pair(sa + a_bytes, &tmB, lbar, kb * BLOCK_R, hh * n_each + (int) rank * n_half);
            }
        }
    } else if (warp == MMA_WARP) {
        if (rank == 0 && elect_one()) {
            const uint32_t idesc = make_idesc(2 * BLOCK_M, n_each, 0, 0);
            uint32_t it = 0;
            for (uint32_t u = 0; u < n_units; ++u) {
                const uint32_t buf = u & 1u, use = u >> 1;
                mbar_wait(&bars->tmem_empty[buf], (use & 1u) ^ 1u);     // the unit that used this buffer has been read out
                tc_fence_after();
                TRACE_FINE(256 + 4 * (int) u);
                const uint32_t acc = tmem_base + buf * BUF_COLS;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = (int) (it % (uint32_t) nstages);
                    const uint32_t ph = (it / (uint32_t) nstages) & 1u;
                    mbar_wait(&bars->full[s], ph);
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(smem + (size_t) s * stage_bytes);
                    const uint32_t b_base = a_base + a_bytes;
#pragma unroll
                    for (int k = 0; k < BLOCK_R / UMMA_K; ++k) {
                        const uint64_t adesc = make_desc(a_base + k * UMMA_K * 4, 16, 8 * BLOCK_R * 4, LAYOUT_SW128);
                        const uint64_t bdesc = make_desc(b_base + k * UMMA_K * 4, 16, 8 * BLOCK_R * 4, LAYOUT_SW128);
                        if (!(ep.dbg & 2)) umma_tf32_pair(acc, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit_pair(&bars->empty[s]);     // frees the stage in BOTH CTAs
                }
                umma_commit_pair(&bars->tmem_full[buf]);   // this unit's accumulator is complete in both CTAs
                TRACE_FINE(256 + 4 * (int) u + 1);
            }
            TRACE(2);
        }
    } else {
        epilogue_prepare<EW>(0, M, N, ep, sbias);
        const uint32_t empty0 = mapa_shared(smem_u32(&bars->tmem_empty[0]), 0);
        const uint32_t empty1 = mapa_shared(smem_u32(&bars->tmem_empty[1]), 0);
        for (uint32_t u = 0; u < n_units; ++u) {
            const int tile = cid + (int) (u / (uint32_t) nmma) * ncl, hh = (int) (u % (uint32_t) nmma);
            const int m0 = (tile * 2 + (int) rank) * BLOCK_M;
            const uint32_t buf = u & 1u, use = u >> 1;
            const int n0 = hh * n_each;
            const int Nu = (N - n0) < n_each ? (N - n0) : n_each;       // columns of this unit
            EpiParams epu = ep;
            if (epu.mask) epu.mask += n0;
            epu.col0 = n0;
            epu.bits_c0 = hh * ((n_each + 31) / 32);
            mbar_wait(&bars->tmem_full[buf], use & 1u);
            tc_fence_after();
            if (threadIdx.x == 0) TRACE_FINE(256 + 4 * (int) u + 2);
            if (Nu > 0 && !(ep.dbg & 64))
                epilogue_dispatch<EW, DIRECT>(tmem_base + buf * BUF_COLS, warp, lane, m0, M, Nu, C + n0, ldc, epu, estage,
                                      smem_u32(sbias) + (uint32_t) (n0 * 4));
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(buf ? empty1 : empty0);
            if (threadIdx.x == 0) TRACE_FINE(256 + 4 * (int) u + 3);
        }
        if (threadIdx.x == 0) TRACE(4);
    }
    tc_fence_before();
    cluster_sync_all();                                   // neither CTA frees TMEM / exits while its peer still runs
    if (threadIdx.x == 0) TRACE(5);
    if (warp == TMA_WARP) tmem_dealloc_pair(tmem_base, TMEM_COLS);
    if (threadIdx.x == TMA_WARP * 32) TRACE(6);
}

// ---- C_split[M,N] = A[R,M]^T * B[R,N] over this CTA's slice of R ; MN-major tf32 tiles (32B-atom swizzle)
constexpr int NT_EW = 8;
__global__ void __launch_bounds__(nthreads(NT_EW), 1)
gemm_nt_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N, int R,
               int r_per_split, int n1, int n2, int nstages, uint32_t tmem_cols, float* __restrict__ Cws, int tma_lanes, int dbg) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int BOX_BYTES = 32 * BLOCK_R * 4;             // 32 MN-elements x 32 reduction rows
    const int nbx = (n1 + n2) / 32;                          // MN blocks of the B tile
    const int a_bytes = (BLOCK_M / 32) * BOX_BYTES;
    const int stage_bytes = a_bytes + nbx * BOX_BYTES;
    const int m0 = blockIdx.x * BLOCK_M;
    const int r_lo = blockIdx.y * r_per_split;
    const int r_hi = (r_lo + r_per_split < R) ? r_lo + r_per_split : R;
    const int num_kb = r_hi > r_lo ? (r_hi - r_lo + BLOCK_R - 1) / BLOCK_R : 0;
    constexpr int EW = NT_EW, TMA_WARP = EW, MMA_WARP = EW + 1;
    const size_t pipe_bytes = pipe_region((size_t) nstages * stage_bytes, EW);
    PipeBars* bars = setup_cta(smem, pipe_bytes, nstages, tmem_cols, warp, TMA_WARP);
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == TMA_WARP) {
        // a k-block is BLOCK_M / 32 + nbx (up to 20) boxes of 32 x 32 floats: one box per lane, so the per-thread TMA issue
        // time (see the tn kernel's producer) is paid once per k-block instead of once per box
        const int nbox = BLOCK_M / 32 + nbx;
        if (tma_lanes > 1) {
            const unsigned active = nbox >= 32 ? 0xffffffffu : ((1u << nbox) - 1u);
            if (lane < nbox) {
                if (lane == 0) {
                    prefetch_tmap(&tmA);
                    prefetch_tmap(&tmB);
                }
                for (int kb = 0; kb < num_kb; ++kb) {
                    const int s = kb % nstages;
                    const uint32_t ph = (uint32_t) (kb / nstages) & 1u;
                    mbar_wait(&bars->empty[s], ph ^ 1u);
                    uint8_t* sa = smem + (size_t) s * stage_bytes;
                    if (lane == 0) mbar_expect_tx(&bars->full[s], (dbg & 4) ? 0u : (uint32_t) stage_bytes);
                    __syncwarp(active);
                    if (dbg & 4) continue;                  // timing experiment: no loads
                    const int r0 = r_lo + kb * BLOCK_R;
                    if (lane < BLOCK_M / 32) tma_load_2d(sa + lane * BOX_BYTES, &tmA, &bars->full[s], m0 + lane * 32, r0);
                    else tma_load_2d(sa + lane * BOX_BYTES, &tmB, &bars->full[s], (lane - BLOCK_M / 32) * 32, r0);
                }
            }
        } else if (elect_one()) {
            prefetch_tmap(&tmA);
            prefetch_tmap(&tmB);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % nstages;
                const uint32_t ph = (uint32_t) (kb / nstages) & 1u;
                mbar_wait(&bars->empty[s], ph ^ 1u);
                uint8_t* sa = smem + (size_t) s * stage_bytes;
                mbar_expect_tx(&bars->full[s], (uint32_t) stage_bytes);
                const int r0 = r_lo + kb * BLOCK_R;
                for (int j = 0; j < BLOCK_M / 32; ++j) tma_load_2d(sa + j * BOX_BYTES, &tmA, &bars->full[s], m0 + j * 32, r0);
                for (int j = 0; j < nbx; ++j) tma_load_2d(sa + a_bytes + j * BOX_BYTES, &tmB, &bars->full[s], j * 32, r0);
            }
        }
    } else if (warp == MMA_WARP) {
        if (elect_one()) {
            const uint32_t idesc1 = make_idesc(BLOCK_M, n1, 1, 1);
            const uint32_t idesc2 = make_idesc(BLOCK_M, n2 > 0 ? n2 : 16, 1, 1);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % nstages;
                const uint32_t ph = (uint32_t) (kb / nstages) & 1u;
                mbar_wait(&bars->full[s], ph);
                tc_fence_after();
                const uint32_t a_base = smem_u32(smem + (size_t) s * stage_bytes);
                const uint32_t b_base = a_base + a_bytes;
#pragma unroll
                for (int k = 0; k < BLOCK_R / UMMA_K; ++k) {
                    // 8 reduction rows = two 4-row swizzle atoms (SBO = 512 B); MN blocks of 32 are LBO apart
                    const uint32_t koff = k * UMMA_K * 128;
                    const uint64_t adesc = make_desc(a_base + koff, BOX_BYTES, 512, LAYOUT_SW128_BASE32B);
                    const uint64_t bdesc1 = make_desc(b_base + koff, BOX_BYTES, 512, LAYOUT_SW128_BASE32B);
                    if (dbg & 2) continue;                  // timing experiment: no MMAs
                    umma_tf32(tmem_base, adesc, bdesc1, idesc1, (kb | k) != 0 ? 1u : 0u);
                    if (n2 > 0) {
                        const uint64_t bdesc2 = make_desc(b_base + (n1 / 32) * BOX_BYTES + koff, BOX_BYTES, 512, LAYOUT_SW128_BASE32B);
                        umma_tf32(tmem_base + (uint32_t) n1, adesc, bdesc2, idesc2, (kb | k) != 0 ? 1u : 0u);
                    }
                }
                umma_commit(&bars->empty[s]);
            }
            umma_commit(&bars->tmem_full);
        }
    } else {
        float* Cz = Cws + (size_t) blockIdx.y * M * N;
        if (num_kb > 0) {
            mbar_wait(&bars->tmem_full, 0);
            tc_fence_after();
            EpiParams ep{nullptr, 0, 1.f, 0, nullptr, 0, 1.f, 0, 0, 0, 0, nullptr, nullptr, 0, 1, nullptr, nullptr, 0, 0};
            if ((N & 3) == 0 && (reinterpret_cast<uintptr_t>(Cz) & 15) == 0)       // partial sums are NOT rounded to tf32
                epilogue_rows_fast<false, false, 0, false, EW, false, true>(tmem_base, warp, lane, m0, M, N, Cz, N, ep, reinterpret_cast<float*>(smem), 0u);
            else
                epilogue_rows<false, false, 0, EW>(tmem_base, warp, lane, m0, M, N, Cz, N, ep, reinterpret_cast<float*>(smem), 0u);
        } else {                                   // empty slice: contribute zeros
            const int row = m0 + (warp & 3) * 32 + lane;
            if (row < M) for (int c = (warp >> 2); c < N; c += EW / 4) Cz[(size_t) row * N + c] = 0.f;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == TMA_WARP) tmem_dealloc(tmem_base, tmem_cols);
}

// ---- the same product on CTA pairs (cta_group::2): ONE M = 256 MMA spans the two CTAs of a cluster, CTA rank r stages
// rows [(2p + r) * 128, +128) of A^T (its own 128 output rows) and HALF of every B tile.  Why: the one-CTA kernel above is
// bound by what the SMs pull out of L2 (ncu r02k: 55.5 us, tensor pipe 52 %; every one of the 4 m-tiles streams all of dZ,
// 525 MB of shared-memory fill per GEMM at the ~9.3 TB/s the L2 fabric delivers, profiles/r02j_probe_loads_vs_stores.txt).
// Sharing B between the two SMs of a pair takes the fill to 315 MB.  N is covered by up to two MMAs per k-step (n1 + n2,
// each a multiple of 64 so that a CTA's half is whole 32-column boxes); columns past N are zero-filled by TMA.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(nthreads(NT_EW), 1)
gemm_nt_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int M, int N, int R,
                    int r_per_split, int n1, int n2, int nstages, float* __restrict__ Cws, int dbg) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    constexpr int BOX_BYTES = 32 * BLOCK_R * 4;             // 32 MN-elements x 32 reduction rows
    const int nb1 = n1 / 64, nb2 = n2 / 64;                  // boxes of THIS CTA's half of each MMA's B tile
    const int nbx = nb1 + nb2;
    const int a_bytes = (BLOCK_M / 32) * BOX_BYTES;
    const int stage_bytes = a_bytes + nbx * BOX_BYTES;
    const int m0 = (int) blockIdx.x * BLOCK_M;               // blockIdx.x = 2 * pair + rank
    const int r_lo = blockIdx.y * r_per_split;
    const int r_hi = (r_lo + r_per_split < R) ? r_lo + r_per_split : R;
    const int num_kb = r_hi > r_lo ? (r_hi - r_lo + BLOCK_R - 1) / BLOCK_R : 0;
    constexpr int EW = NT_EW, TMA_WARP = EW, MMA_WARP = EW + 1;
    constexpr uint32_t TMEM_COLS = 512;
    const size_t pipe_bytes = pipe_region((size_t) nstages * stage_bytes, EW);
    PipeBars* bars = reinterpret_cast<PipeBars*>(smem + pipe_bytes);
    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        mbar_init(&bars->tmem_full, 1);
        fence_barrier_init();
    }
    cluster_sync_all();                                      // both CTAs' barriers exist before anyone signals them
    if (warp == TMA_WARP) tmem_alloc_pair(&bars->tmem_base, TMEM_COLS);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    if (warp == TMA_WARP) {
        // one box per lane (see the tn kernel's producer): 4 boxes of A, nbx of this CTA's B halves
        const int nbox = BLOCK_M / 32 + nbx;
        const unsigned active = (1u << nbox) - 1u;           // nbox <= 4 + 7
        if (lane < nbox) {
            if (lane == 0) {
                prefetch_tmap(&tmA);
                prefetch_tmap(&tmB);
            }
            // column of B this lane's box starts at: MMA 1 covers [0, n1), this CTA's half is [rank * n1 / 2, +n1 / 2);
            // MMA 2 covers [n1, n1 + n2) likewise
            int bcol = 0;
            if (lane >= BLOCK_M / 32) {
                const int j = lane - BLOCK_M / 32;
                bcol = j < nb1 ? (int) rank * (n1 / 2) + j * 32 : n1 + (int) rank * (n2 / 2) + (j - nb1) * 32;
            }
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % nstages;
                const uint32_t ph = (uint32_t) (kb / nstages) & 1u;
                mbar_wait(&bars->empty[s], ph ^ 1u);           // my copy of the stage is free (multicast commit)
                uint8_t* sa = smem + (size_t) s * stage_bytes;
                if (rank == 0 && lane == 0) mbar_expect_tx(&bars->full[s], (dbg & 4) ? 0u : (uint32_t) (2 * stage_bytes));
                __syncwarp(active);
                if (dbg & 4) continue;
                const uint32_t lbar = mapa_shared(smem_u32(&bars->full[s]), 0);
                const int r0 = r_lo + kb * BLOCK_R;
                if (lane < BLOCK_M / 32) tma_load_2d_pair(sa + lane * BOX_BYTES, &tmA, lbar, m0 + lane * 32, r0);
                else tma_load_2d_pair(sa + lane * BOX_BYTES, &tmB, lbar, bcol, r0);
            }
        }
    } else if (warp == MMA_WARP) {
        if (rank == 0 && elect_one()) {
            const uint32_t idesc1 = make_idesc(2 * BLOCK_M, n1, 1, 1);
            const uint32_t idesc2 = make_idesc(2 * BLOCK_M, n2 > 0 ? n2 : 64, 1, 1);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % nstages;
                const uint32_t ph = (uint32_t) (kb / nstages) & 1u;
                mbar_wait(&bars->full[s], ph);
                tc_fence_after();
                const uint32_t a_base = smem_u32(smem + (size_t) s * stage_bytes);
                const uint32_t b_base = a_base + a_bytes;
#pragma unroll
                for (int k = 0; k < BLOCK_R / UMMA_K; ++k) {
                    if (dbg & 2) continue;
                    const uint32_t koff = k * UMMA_K * 128;
                    const uint64_t adesc = make_desc(a_base + koff, BOX_BYTES, 512, LAYOUT_SW128_BASE32B);
                    const uint64_t bdesc1 = make_desc(b_base + koff, BOX_BYTES, 512, LAYOUT_SW128_BASE32B);
                    umma_tf32_pair(tmem_base, adesc, bdesc1, idesc1, (kb | k) != 0 ? 1u : 0u);
                    if (n2 > 0) {
                        const uint64_t bdesc2 = make_desc(b_base + nb1 * BOX_BYTES + koff, BOX_BYTES, 512, LAYOUT_SW128_BASE32B);
                        umma_tf32_pair(tmem_base + (uint32_t) n1, adesc, bdesc2, idesc2, (kb | k) != 0 ? 1u : 0u);
                    }
                }
                umma_commit_pair(&bars->empty[s]);     // frees the stage in BOTH CTAs
            }
            umma_commit_pair(&bars->tmem_full);        // accumulators complete in both CTAs
        }
    } else {
        float* Cz = Cws + (size_t) blockIdx.y * M * N;
        if (num_kb > 0) {
            mbar_wait(&bars->tmem_full, 0);
            tc_fence_after();
            EpiParams ep{nullptr, 0, 1.f, 0, nullptr, 0, 1.f, 0, 0, 0, 0, nullptr, nullptr, 0, 1, nullptr, nullptr, 0, 0};
            if (m0 < M) {
                if ((N & 3) == 0 && (reinterpret_cast<uintptr_t>(Cz) & 15) == 0)   // partial sums are NOT rounded to tf32
                    epilogue_rows_fast<false, false, 0, false, EW, false, true>(tmem_base, warp, lane, m0, M, N, Cz, N, ep, reinterpret_cast<float*>(smem), 0u);
                else
                    epilogue_rows<false, false, 0, EW>(tmem_base, warp, lane, m0, M, N, Cz, N, ep, reinterpret_cast<float*>(smem), 0u);
            }
        } else {                                   // empty slice: contribute zeros
            const int row = m0 + (warp & 3) * 32 + lane;
            if (row < M) for (int c = (warp >> 2); c < N; c += EW / 4) Cz[(size_t) row * N + c] = 0.f;
        }
    }
    tc_fence_before();
    cluster_sync_all();                                   // neither CTA frees TMEM / exits while its peer still runs
    if (warp == TMA_WARP) tmem_dealloc_pair(tmem_base, TMEM_COLS);
}

}  // namespace

// ------------------------------------------------------------------------------------ host side
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// 2-D fp32 tensor [rows, cols] with row pitch ld floats; box = {box_cols, box_rows}; OOB reads give zeros
int make_map(ctr_handle* h, CUtensorMap* map, const float* base, int64_t rows, int64_t cols, int64_t ld, int box_cols,
             int box_rows, CUtensorMapSwizzle swz) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return ctr_fail(h, CTR_E_UNSUPPORTED, "cuTensorMapEncodeTiled not available");
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld % 4) != 0)
        return ctr_fail(h, CTR_E_INVALID, "TF32 path: operand must be 16-byte aligned with a row pitch multiple of 4 floats (ld=%lld)",
                        (long long) ld);
    cuuint64_t dims[2] = {(cuuint64_t) cols, (cuuint64_t) rows};
    cuuint64_t strides[1] = {(cuuint64_t) ld * 4};
    cuuint32_t box[2] = {(cuuint32_t) box_cols, (cuuint32_t) box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return ctr_fail(h, CTR_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int) r);
    return CTR_OK;
}

uint32_t tmem_cols_for(int nb) {
    uint32_t c = 32;
    while ((int) c < nb) c <<= 1;
    return c;
}

int pick_stages(int stage_bytes, int cap = 4) {
    int s = (SMEM_LIMIT - 2048 - SBIAS_BYTES) / stage_bytes;
    return s > cap ? cap : s;
}

unsigned long long* g_trace_buf = nullptr;
unsigned long long* gemm_trace_buf();
int gemm_dbg() {
    static int d = -1;
    if (d < 0) {
        const char* e = getenv("CTR_GEMM_DBG");
        d = e ? atoi(e) : 0;
    }
    return d;
}

unsigned long long* gemm_trace_buf() {
    if (!(gemm_dbg() & 8)) return nullptr;
    if (!g_trace_buf) {
        cudaMalloc(&g_trace_buf, (4096 * 8 + 512) * sizeof(unsigned long long));
        cudaMemset(g_trace_buf, 0, (4096 * 8 + 512) * sizeof(unsigned long long));
    }
    return g_trace_buf;
}

// CTR_GEMM_TMA_LANES = lanes of the producer warp that issue TMA loads in the persistent tn kernel (k-blocks dealt
// round-robin) and in the nt kernel (> 1: one box per lane).  Default 4; 1 = one elected thread issues everything (the
// round-1 kernels, kept for A/B).
int tma_lanes() {
    static int n = -1;
    if (n < 0) {
        const char* e = getenv("CTR_GEMM_TMA_LANES");
        n = e ? atoi(e) : 4;
        if (n < 1) n = 1;
        if (n > 8) n = 8;
    }
    return n;
}

// CTR_GEMM_MODE = pp (default: persistent CTA pairs, double-buffered accumulator) | split2 (two co-resident half-N CTAs
// per SM) | pair (one cta_group::2 tile per cluster) | single (one 128-row tile per CTA)
int gemm_mode() {
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("CTR_GEMM_MODE");
        mode = 3;
        if (e && e[0] == 's' && e[1] == 'p') mode = 0;
        if (e && e[0] == 'p' && e[1] == 'a') mode = 1;
        if (e && e[0] == 's' && e[1] == 'i') mode = 2;
    }
    return mode;
}

int launch_tn(ctr_handle* h, cudaStream_t st, int M, int N, int R, const float* A, int64_t lda, const float* B,
              int64_t ldb, float* C, int64_t ldc, EpiParams ep) {
    CTR_REQUIRE(h, N <= MAX_N, "TF32 path: N=%d > %d not supported", N, MAX_N);
    const int nmma = N > 256 ? 2 : 1;
    const int NB = nmma == 1 ? (int) ctr_align(N, 16) : (int) ctr_align(N, 32);
    const int n_each = NB / nmma;
    // sign-bit slots per 32-row block: the units / slabs of every kernel below are n_each columns wide and start their
    // sub-tile count at bits_c0 = index * ceil(n_each / 32); a CTA that covers all NB columns in one sweep (single, pair)
    // counts ceil(NB / 32) <= nmma * ceil(n_each / 32) sub-tiles from 0
    ep.bits_chunks = nmma * ((n_each + 31) / 32);
    // output columns [N, align4(N)) exist in C when ldc allows (pad columns of the activations / dX0): storing the zeros
    // the accumulator holds there keeps every store a full 128-bit one (fast epilogue).  Not with a bias / mask, whose
    // vectors end at N.
    const int Ne = ((N & 3) && ldc >= ctr_align(N, 4) && !ep.bias && !ep.mask) ? (int) ctr_align(N, 4) : N;
    if (gemm_mode() == 3 && M > BLOCK_M) {
        // persistent CTA pairs, double-buffered accumulator (see gemm_tn_pair_persistent_kernel)
        const int n_half = n_each / 2;
        const int stage_bytes = BLOCK_M * BLOCK_R * 4 + n_half * BLOCK_R * 4;
        int nstages = (int) ((SMEM_LIMIT - 1024 - SBIAS_BYTES - sizeof(PairBars) - 8 * EPI_TILE_FLOATS * 4) / stage_bytes);
        if (nstages > 6) nstages = 6;
        CTR_REQUIRE(h, nstages >= 2, "TF32 path: tile does not fit shared memory");
        CUtensorMap tmA, tmB;
        int rc = make_map(h, &tmA, A, M, R, lda, BLOCK_R, BLOCK_M, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
        rc = make_map(h, &tmB, B, N, R, ldb, BLOCK_R, n_half, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
        const size_t smem = pair_persistent_smem((size_t) nstages * stage_bytes, 8);
        static bool attr_set3 = false;
        if (!attr_set3) {
            CTR_CUDA(h, cudaFuncSetAttribute(gemm_tn_pair_persistent_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
            attr_set3 = true;
        }
        const int n_tiles = (int) ctr_cdiv(M, 2 * BLOCK_M);
        const int max_cl = h->sm_count / 2;
        const unsigned ctas = 2u * (unsigned) (n_tiles < max_cl ? n_tiles : max_cl);
        if (gemm_dbg() & 16) {      // experiment: transposition-free epilogue on 256-bit stores (separate instantiation, so
                                    // the default kernel's code and register allocation stay exactly as measured)
            static bool attr_direct = false;
            if (!attr_direct) {
                CTR_CUDA(h, cudaFuncSetAttribute(gemm_tn_pair_persistent_kernel<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
                attr_direct = true;
            }
            gemm_tn_pair_persistent_kernel<8, true><<<ctas, nthreads(8), smem, st>>>(tmA, tmB, M, Ne, R, n_each, nmma, nstages, C, ldc, ep);
        } else
        gemm_tn_pair_persistent_kernel<8><<<ctas, nthreads(8), smem, st>>>(tmA, tmB, M, Ne, R, n_each, nmma, nstages, C, ldc, ep);
        CTR_CHECK_LAUNCH(h, "gemm_tn_pair_persistent_kernel");
        return CTR_OK;
    }
    if (gemm_mode() == 1 && M > BLOCK_M) {
        // CTA pairs: each CTA stages 128 rows of A and half of B per k-block
        const int n_half = n_each / 2;
        const int stage_bytes = BLOCK_M * BLOCK_R * 4 + nmma * n_half * BLOCK_R * 4;
        const int nstages = pick_stages(stage_bytes, 6);
        CTR_REQUIRE(h, nstages >= 2, "TF32 path: tile does not fit shared memory");
        CUtensorMap tmA, tmB;
        int rc = make_map(h, &tmA, A, M, R, lda, BLOCK_R, BLOCK_M, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
        rc = make_map(h, &tmB, B, N, R, ldb, BLOCK_R, n_half, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
        const size_t smem = smem_total((size_t) nstages * stage_bytes, 8);
        static bool attr_set2 = false;
        if (!attr_set2) {
            CTR_CUDA(h, cudaFuncSetAttribute(gemm_tn_pair_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
            attr_set2 = true;
        }
        const unsigned ctas = (unsigned) ctr_align(ctr_cdiv(M, BLOCK_M), 2);
        gemm_tn_pair_kernel<8><<<ctas, nthreads(8), smem, st>>>(tmA, tmB, M, Ne, R, n_each, nmma, nstages, tmem_cols_for(NB), C,
                                                             ldc, ep);
        CTR_CHECK_LAUNCH(h, "gemm_tn_pair_kernel");
        return CTR_OK;
    }
    CUtensorMap tmA, tmB;
    int rc = make_map(h, &tmA, A, M, R, lda, BLOCK_R, BLOCK_M, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    rc = make_map(h, &tmB, B, N, R, ldb, BLOCK_R, n_each, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        CTR_CUDA(h, cudaFuncSetAttribute(gemm_tn_kernel<1, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
        CTR_CUDA(h, cudaFuncSetAttribute(gemm_tn_kernel<2, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT / 2));
        attr_set = true;
    }
    // split2: each CTA takes ONE MMA slab of n_each (<= 256) columns; two such CTAs fit an SM (256 TMEM columns and
    // 2-3 pipeline stages each), which overlaps epilogue / fill / MMA of different tiles without a persistent scheduler
    const int slab_stage = BLOCK_M * BLOCK_R * 4 + n_each * BLOCK_R * 4;
    const int slab_budget = SMEM_LIMIT / 2 - 2048 - 1024 - SBIAS_BYTES;
    if (gemm_mode() == 0 && slab_budget / slab_stage >= 2) {
        int nstages = slab_budget / slab_stage;
        if (nstages > 4) nstages = 4;
        const size_t smem = smem_total((size_t) nstages * slab_stage, 8);
        const unsigned grid = (unsigned) (ctr_cdiv(M, BLOCK_M) * nmma);
        static int stagger = -1;
        static unsigned* slots = nullptr;
        if (stagger < 0) {
            const char* e = getenv("CTR_GEMM_STAGGER_NS");
            stagger = e ? atoi(e) : 0;
            if (stagger > 0) CTR_CUDA(h, cudaMalloc(&slots, 1024 * sizeof(unsigned)));
        }
        EpiParams ep2 = ep;
        if (stagger > 0) {
            CTR_CUDA(h, cudaMemsetAsync(slots, 0, 1024 * sizeof(unsigned), st));
            ep2.sm_slots = slots;
            ep2.stagger_ns = (unsigned) stagger;
        }
        gemm_tn_kernel<2, 8><<<grid, nthreads(8), smem, st>>>(tmA, tmB, M, Ne, R, n_each, 1, nmma, nstages, tmem_cols_for(n_each), C,
                                                           ldc, ep2);
        CTR_CHECK_LAUNCH(h, "gemm_tn_kernel<2>");
        return CTR_OK;
    }
    const int stage_bytes = BLOCK_M * BLOCK_R * 4 + NB * BLOCK_R * 4;
    const int nstages = pick_stages(stage_bytes);
    CTR_REQUIRE(h, nstages >= 2, "TF32 path: tile does not fit shared memory");
    const size_t smem = smem_total((size_t) nstages * stage_bytes, 8);
    gemm_tn_kernel<1, 8><<<(unsigned) ctr_cdiv(M, BLOCK_M), nthreads(8), smem, st>>>(tmA, tmB, M, Ne, R, n_each, nmma, 1, nstages,
                                                                                 tmem_cols_for(NB), C, ldc, ep);
    CTR_CHECK_LAUNCH(h, "gemm_tn_kernel");
    return CTR_OK;
}

struct NtPlan { int n1, n2, nstages, splits, r_per_split, mtiles; };

NtPlan plan_nt(ctr_handle* h, int M, int N, int R) {
    NtPlan p;
    const int NB = (int) ctr_align(N, 32);
    p.n1 = NB > 256 ? 256 : NB;
    p.n2 = NB - p.n1;
    const int stage_bytes = (BLOCK_M / 32 + NB / 32) * 32 * BLOCK_R * 4;
    p.nstages = pick_stages(stage_bytes);
    p.mtiles = (int) ctr_cdiv(M, BLOCK_M);
    int want = h->sm_count / p.mtiles;
    if (want < 1) want = 1;
    int64_t rps = ctr_align(ctr_cdiv(R, want), BLOCK_R);
    if (rps < 4 * BLOCK_R) rps = 4 * BLOCK_R;
    p.r_per_split = (int) rps;
    p.splits = (int) ctr_cdiv(R, rps);
    return p;
}

// the CTA-pair form (gemm_nt_pair_kernel): pairs of m-tiles, B columns in MMAs of n1 + n2 (multiples of 64)
NtPlan plan_nt_pair(ctr_handle* h, int M, int N, int R) {
    NtPlan p;
    const int NB = (int) ctr_align(N, 64);
    p.n1 = NB > 256 ? 256 : NB;
    p.n2 = NB - p.n1;
    const int stage_bytes = (BLOCK_M / 32 + NB / 64) * 32 * BLOCK_R * 4;
    p.nstages = pick_stages(stage_bytes);
    p.mtiles = 2 * (int) ctr_cdiv(ctr_cdiv(M, BLOCK_M), 2);        // CTAs along x: whole pairs
    int want = h->sm_count / p.mtiles;
    if (want < 1) want = 1;
    int64_t rps = ctr_align(ctr_cdiv(R, want), BLOCK_R);
    if (rps < 4 * BLOCK_R) rps = 4 * BLOCK_R;
    p.r_per_split = (int) rps;
    p.splits = (int) ctr_cdiv(R, rps);
    return p;
}

// CTR_WGRAD = pair (default) | single: which weight-gradient kernel runs (A/B)
bool wgrad_pair() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("CTR_WGRAD");
        v = (e && e[0] == 's') ? 0 : 1;
    }
    return v == 1;
}

// out[i] = sum over the splits; elements past n_main (the extra row produced by a ones column of X) go to `tail`
__global__ void __launch_bounds__(256) tc_splitk_reduce(const float* __restrict__ ws, int splits, int64_t n, int64_t n_main,
                                                        float* __restrict__ out, float* __restrict__ tail) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        float a0 = 0.f, a1 = 0.f;
        int z = 0;
        for (; z + 2 <= splits; z += 2) {
            a0 += ws[(int64_t) z * n + i];
            a1 += ws[(int64_t) (z + 1) * n + i];
        }
        if (z < splits) a0 += ws[(int64_t) z * n + i];
        const float a = a0 + a1;
        if (i < n_main) out[i] = a;
        else tail[i - n_main] = a;
    }
}

}  // namespace

int ctr_tc_mlp_fwd(ctr_handle* h, cudaStream_t st, int B, int K, int N, const float* X, int64_t ldx, const float* Wt,
                   const float* bias, int relu, float keep, uint64_t seed, float* Y, int64_t ldy, uint32_t* relu_bits) {
    // Wt is the K-major copy [N, ldwt] with ldwt = K rounded up to 4 floats (TMA needs 16-byte row pitches)
    EpiParams ep{bias, relu, keep, seed, nullptr, 0, 1.f, 1, 0, N, gemm_dbg(), gemm_trace_buf(), nullptr, 0, tma_lanes(),
                 relu_bits, nullptr, 0, 0};
    return launch_tn(h, st, B, N, K, X, ldx, Wt, ctr_align(K, 4), Y, ldy, ep);
}

// bytes of the sign-bit buffer of a [B, N] activation: 4 blocks of 32 rows per 128-row tile, one word per lane and slot
int64_t ctr_tc_relu_bits_bytes(int B, int N) {
    const int nmma = N > 256 ? 2 : 1;
    const int NB = nmma == 1 ? (int) ctr_align(N, 16) : (int) ctr_align(N, 32);
    const int64_t slots = (int64_t) nmma * ((NB / nmma + 31) / 32);
    return ctr_align(ctr_cdiv(B, BLOCK_M) * 4 * slots * 32 * 4, 256);
}

int ctr_tc_mlp_bwd_data(ctr_handle* h, cudaStream_t st, int B, int K, int N, const float* dZ, int64_t lddz,
                        const float* W, const float* mask, int64_t ldmask, float mask_scale, float* dX, int64_t lddx,
                        const uint32_t* relu_bits) {
    // dX[B,K] = dZ[B,N] * W[K,N]^T : W is already reduction-major for this product
    EpiParams ep{nullptr, 0, 1.f, 0, relu_bits ? nullptr : mask, ldmask, mask_scale, 1, 0, K, gemm_dbg(), gemm_trace_buf(), nullptr, 0,
                 tma_lanes(), nullptr, relu_bits, 0, 0};
    return launch_tn(h, st, B, K, N, dZ, lddz, W, N, dX, lddx, ep);
}

int64_t ctr_tc_bwd_weight_ws_bytes(ctr_handle* h, int B, int K, int N) {
    NtPlan p = plan_nt(h, K + 1, N, B);       // + 1: room for the dbias row of the ones-column form
    NtPlan q = plan_nt_pair(h, K + 1, N, B);
    const int splits = p.splits > q.splits ? p.splits : q.splits;     // either kernel may run
    return ctr_align((int64_t) splits * (K + 1) * N * 4, 256);
}

// dbias != nullptr: X carries a column of ones at index K_w (ldx > K_w), so the GEMM is run with K_w + 1 rows and
// its last row, 1^T dZ, is the bias gradient: no separate pass over dZ.
int ctr_tc_mlp_bwd_weight(ctr_handle* h, cudaStream_t st, int B, int K_w, int N, const float* X, int64_t ldx,
                          const float* dZ, int64_t lddz, float* dW, float* dbias, void* ws, int64_t ws_bytes) {
    CTR_REQUIRE(h, N <= MAX_N, "TF32 path: N=%d > %d not supported", N, MAX_N);
    const int K = K_w + (dbias ? 1 : 0);
    CTR_REQUIRE(h, ldx >= K, "TF32 wgrad: ones-column form needs ldx > K");
    // CTA pairs only where there are at least two real m-tiles to pair (with one, the second CTA of the pair would idle and the
    // one-CTA kernel has twice the batch splits)
    const bool pair = wgrad_pair() && B >= 4 * BLOCK_R && K > BLOCK_M;
    NtPlan p = pair ? plan_nt_pair(h, K, N, B) : plan_nt(h, K, N, B);
    CTR_REQUIRE(h, p.nstages >= 2, "TF32 wgrad: tile does not fit shared memory");
    if (ws_bytes < (int64_t) p.splits * K * N * 4) return ctr_fail(h, CTR_E_WORKSPACE, "TF32 wgrad: ws too small");
    CUtensorMap tmA, tmB;
    int rc = make_map(h, &tmA, X, B, K, ldx, 32, BLOCK_R, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
    if (rc) return rc;
    rc = make_map(h, &tmB, dZ, B, N, lddz, 32, BLOCK_R, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
    if (rc) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        CTR_CUDA(h, cudaFuncSetAttribute(gemm_nt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
        CTR_CUDA(h, cudaFuncSetAttribute(gemm_nt_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
        attr_set = true;
    }
    dim3 grid((unsigned) p.mtiles, (unsigned) p.splits);
    if (pair) {
        const int stage_bytes = (BLOCK_M / 32 + (p.n1 + p.n2) / 64) * 32 * BLOCK_R * 4;
        const size_t smem = smem_total((size_t) p.nstages * stage_bytes, NT_EW);
        gemm_nt_pair_kernel<<<grid, nthreads(NT_EW), smem, st>>>(tmA, tmB, K, N, B, p.r_per_split, p.n1, p.n2, p.nstages, (float*) ws,
                                                             gemm_dbg());
        CTR_CHECK_LAUNCH(h, "gemm_nt_pair_kernel");
    } else {
        const int stage_bytes = (BLOCK_M / 32 + (p.n1 + p.n2) / 32) * 32 * BLOCK_R * 4;
        const size_t smem = smem_total((size_t) p.nstages * stage_bytes, NT_EW);
        gemm_nt_kernel<<<grid, nthreads(NT_EW), smem, st>>>(tmA, tmB, K, N, B, p.r_per_split, p.n1, p.n2, p.nstages,
                                                        tmem_cols_for(p.n1 + p.n2), (float*) ws, tma_lanes(), gemm_dbg());
        CTR_CHECK_LAUNCH(h, "gemm_nt_kernel");
    }
    const int64_t n = (int64_t) K * N;
    tc_splitk_reduce<<<(unsigned) ctr_cdiv(n, 256), 256, 0, st>>>((const float*) ws, p.splits, n, (int64_t) K_w * N, dW, dbias);
    CTR_CHECK_LAUNCH(h, "tc_splitk_reduce");
    return CTR_OK;
}

// timing experiments only: copies the per-CTA phase timestamps of the last traced GEMM (CTR_GEMM_DBG & 8) to the host
extern "C" int ctr_debug_gemm_trace(unsigned long long* host, int n_ctas) {
    if (!g_trace_buf || !host || n_ctas <= 0 || n_ctas > 4096 + 64) return CTR_E_INVALID;
    return cudaMemcpy(host, g_trace_buf, (size_t) n_ctas * 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost) == cudaSuccess
               ? CTR_OK : CTR_E_CUDA;
}
