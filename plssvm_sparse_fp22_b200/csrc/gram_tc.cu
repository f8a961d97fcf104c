// tcgen05 / TMA / TMEM Gram-tile path for dense float data (BASELINE configs[1]: rbf, 65536 x 1024, float).
//
// What it computes (reference semantics: src/plssvm/backends/OpenMP/svm_kernel.cpp:21-47, lower triangle mirrored):
//   for every 128 x 256 tile (I, J) of the lower block triangle of the implicit matrix
//       G    = X_I X_J^T                         (tensor cores, fp32 accumulation in TMEM)
//       t_ij = k(G_ij) + QA_cost - q_i - q_j     (kernel function + the reference's per-entry centring, svm_kernel.cpp:32)
//       Pr[J][i] = sum_j t_ij d_j                (row contribution,   i >= j)
//       Pc[I][j] = sum_i t_ij d_i                (mirrored contribution, i > j)
//   a small deterministic pass (k_tc_reduce) then sums the partial vectors in a fixed order — no atomics.
//
// Precision: X is split once into two fp16 planes, x * s = hi + lo (s = power of two placing max|x| in [2^13, 2^14)), and
//   G = (hi hi^T + hi lo^T + lo hi^T) / s^2 (the dropped lo lo^T term is 2^-22 relative).  Two ways to form the cross terms:
//   MODE_F16X3 : three kind::f16 MMAs per K-slice (22 significant bits per operand);
//   MODE_FP8X  : hi hi^T in kind::f16, the two cross terms as ONE kind::f8f6f4 chain over K-concatenated e4m3 planes
//                [hi8 | lo8] x [lo8 | hi8]^T with hi8 = e4m3(hi 2^-6), lo8 = e4m3(lo 2^6) — fp8 runs at twice the f16 rate, so a
//                K-slice costs 2 instead of 3 pass-equivalents.  The cross terms are 2^-11 of the main term, their 4-bit
//                operands make an error of ~2^-15 of a product: the same size as the fp32 accumulation noise of a 1024-term
//                dot product (DESIGN.md §3.1 has the measured error of both modes).
//
// Structure (one persistent CTA per SM, 320 threads):
//   warp 0 / lane 0 : TMA producer  — cp.async.bulk.tensor 2-D boxes (128 bytes x 128 rows, SWIZZLE_128B) into a ring of four 48 KB slots
//   warp 1 / lane 0 : MMA issuer    — tcgen05.mma.cta_group::1.kind::f16, M=128 N=256 K=16, accumulator double-buffered in TMEM
//   warps 2..9      : epilogue      — (lane quadrant, column half) each; tcgen05.ld 32x32b, exp2 / pow, centring, row sums in registers, column sums by a
//                                     warp-shuffle butterfly + shared memory
#include "internal.cuh"
#include "xchg.cuh"

#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include <algorithm>
#include <cmath>
#include <cstring>

namespace b200 {

namespace tc {

constexpr int BM = 128;       // tile rows   (A operand, "i")
constexpr int BN = 256;       // tile cols   (B operand, "j")
constexpr int BK = 64;        // fp16 elements per K block = 128 bytes = one SWIZZLE_128B row
constexpr int STAGES = 4;                                 // ring of 48 KB slots; a K block uses two: (A_hi, B_hi) then the cross-term operands
constexpr int UMMA_K = 16;
constexpr int A_BYTES = BM * BK * 2;                      // 16 KB per plane
constexpr int B_BYTES = BN * BK * 2;                      // 32 KB per plane
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;            // one A box + two B boxes: 48 KB
constexpr int EPI_WARPS = 8;                              // two per SM sub-partition: one per (TMEM lane quadrant, column half)
constexpr int NUM_THREADS = 64 + EPI_WARPS * 32;          // TMA warp + MMA warp + epilogue warps = 320
constexpr int TMEM_COLS = 512;                            // two 256-column fp32 accumulators

struct smem_layout {
    // operand ring (1024-byte aligned for SWIZZLE_128B): per stage [A_hi | A_lo | B_hi | B_lo]
    static constexpr int OPERANDS = 0;
    static constexpr int COLBUF = STAGES * STAGE_BYTES;              // float4[2][BN]   per-column {c_j, q_j, d_j, -}
    static constexpr int COLRED = COLBUF + 2 * BN * 16;              // float[2][4][BN] per-warp column partial sums
    static constexpr int ROWRED = COLRED + 2 * 4 * BN * 4;           // float[2][BM] row sums of the upper column half
    static constexpr int BARRIERS = ROWRED + 2 * BM * 4;             // uint64: full[S], empty[S], tmem_full[2], tmem_empty[2]
    static constexpr int TMEM_PTR = BARRIERS + (2 * STAGES + 4) * 8;
    static constexpr int TOTAL = TMEM_PTR + 16;
};
constexpr int SMEM_BYTES = smem_layout::TOTAL + 1024;  // + slack for the manual 1024-byte alignment

struct params {
    const int2 *tiles;
    int num_tiles;
    int num_k_blocks;
    const float4 *pk;  // [n_pad] {c_i, q_i, d_i, 0}; rows >= m are all-zero (d = 0 masks them)
    float *Pr;         // [T_c][ldP]
    float *Pc;         // [T_r][ldP]
    int64_t ldP;
    float w;           // rbf: 2 gamma log2(e) / s^2 ; polynomial: gamma / s^2
    float coef0;
    float qa;          // QA_cost
    int degree;
    int kernel;
    float *Kc;          // STORE mode: kernel values of every tile, [num_tiles][BN / 4][BM] float4 (kc_quad_index)
    const float *done;  // CG convergence flag (scal[S_DONE]): a launch queued behind the converged iteration returns at once
};

// ---------------------------------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra WAIT_DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t"
        "}" ::"r"(bar),
        "r"(parity)
        : "memory");
}
// `policy`: L2 eviction priority of the fetched lines (createpolicy): the operand planes are re-read by every tile of a super-tile
// and must outlive the kernel-value tiles the STORE pass streams through L2 (evict_last against the stores' evict_first)
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, uint64_t policy) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
                 "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "l"(policy)
                 : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc_mma_f8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// K-major SWIZZLE_128B shared-memory matrix descriptor (sm_100 "version 1"): rows of 128 bytes, 8-row groups 1024 bytes apart.
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((saddr & 0x3FFFF) >> 4);   // start address, bits [0,14)
    d |= static_cast<uint64_t>(1) << 16;                  // leading byte offset (unused for swizzled K-major), bits [16,30)
    d |= static_cast<uint64_t>(1024 >> 4) << 32;          // stride byte offset between 8-row groups, bits [32,46)
    d |= static_cast<uint64_t>(1) << 46;                  // descriptor version (Blackwell), bits [46,48)
    d |= static_cast<uint64_t>(2) << 61;                  // layout type SWIZZLE_128B, bits [61,64)
    return d;
}
// instruction descriptor: D = f32, A = B = f16, both K-major, N at bits [17,23) (N >> 3), M at bits [24,29) (M >> 4)
__host__ __device__ constexpr uint32_t make_idesc_f16(int M, int N) {
    return (1u << 4) | (0u << 7) | (0u << 10) | (0u << 15) | (0u << 16) | (static_cast<uint32_t>(N >> 3) << 17) |
           (static_cast<uint32_t>(M >> 4) << 24);
}

// ---------------------------------------------------------------------------------------------------------------------
// epilogue math
// ---------------------------------------------------------------------------------------------------------------------
template <int KERNEL>
__device__ __forceinline__ float kernel_from_gram(float g, float ci, float cj, float w, float coef0, int degree) {
    if constexpr (KERNEL == B200SVM_RBF) {
        // exp(-gamma (n_i + n_j - 2 G)) = 2^(c_i + c_j + w G), c = -gamma log2(e) n, w = 2 gamma log2(e) / s^2 ; clamp d2 >= 0
        const float a = fminf(fmaf(w, g, ci) + cj, 0.f);
        return ex2_approx(a);
    } else {
        float base = fmaf(w, g, coef0);
        const bool neg = degree < 0;
        unsigned int u = neg ? static_cast<unsigned int>(-degree) : static_cast<unsigned int>(degree);
        float r = 1.f;
        while (u != 0u) {
            if (u & 1u) { r *= base; }
            base *= base;
            u >>= 1;
        }
        return neg ? 1.f / r : r;
    }
}

// Stored kernel-matrix tile (STORE pass -> k_tile_sweep): quad-column-major, float4 index of (columns 4 cq .. 4 cq + 3, row r).
// Both kernels put 32 consecutive rows on the 32 lanes of a warp, so every global access is 512 contiguous bytes.
__host__ __device__ constexpr int kc_quad_index(int cq, int r) { return cq * BM + r; }

// Sum v[c] over the 32 lanes of the warp for each of the 32 columns c; lane l returns the total of column l.
// Fixed exchange pattern => bit-reproducible.
__device__ __forceinline__ float warp_column_sums(float (&v)[32], int lane) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int k = 0; k < o; ++k) {
            const float send = up ? v[k] : v[k + o];
            const float keep = up ? v[k + o] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return v[0];
}

// FP8X: map_lo / map_b are the e4m3 planes [hi8 | lo8] (A role) and [lo8 | hi8] (B role), 128 bytes per 64-element K block;
// otherwise map_lo is the fp16 lo plane and map_b is unused.  Every K block fills two ring slots of [A 16K | B 32K]:
//   slot 0 = (A_hi, B_hi), slot 1 = (A_lo, B_lo) for F16X3 or (A_8, B_8) for FP8X.
// STORE: the epilogue also writes the kernel values k(x_i, x_j) of the tile to p.Kc (the matvec that assembles the cache needs no sweep)
template <int KERNEL, bool FP8X, bool STORE>
__global__ void __launch_bounds__(NUM_THREADS, 1)
    k_gram_tc(const __grid_constant__ CUtensorMap map_hi, const __grid_constant__ CUtensorMap map_lo, const __grid_constant__ CUtensorMap map_b,
              const params p) {
    extern __shared__ uint8_t smem_raw[];
    if (p.done != nullptr && *p.done != 0.f) { return; }  // uniform over the grid (the flag is only written between launches)
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    const uint32_t bar_full = sbase + smem_layout::BARRIERS;          // [STAGES]
    const uint32_t bar_empty = bar_full + STAGES * 8;                  // [STAGES]
    const uint32_t bar_tfull = bar_empty + STAGES * 8;                 // [2]
    const uint32_t bar_tempty = bar_tfull + 2 * 8;                     // [2]
    uint32_t *tmem_ptr_smem = reinterpret_cast<uint32_t *>(smem + smem_layout::TMEM_PTR);

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar_full + s * 8, 1);
            mbar_init(bar_empty + s * 8, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(bar_tfull + b * 8, 1);
            mbar_init(bar_tempty + b * 8, EPI_WARPS);  // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    const int first_tile = blockIdx.x, tile_step = gridDim.x;

    if (warp == 0) {
        // ============================== TMA producer ==============================
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_hi)) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_lo)) : "memory");
            if constexpr (FP8X) { asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory"); }
            const uint64_t keep = l2_policy_evict_last();
            int stage = 0;
            uint32_t phase = 0;
            for (int t = first_tile; t < p.num_tiles; t += tile_step) {
                const int2 tile = p.tiles[t];
                const int row_a = tile.x * BM, row_b = tile.y * BN;
                for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                    const int kc = kb * BK;
#pragma unroll
                    for (int half = 0; half < 2; ++half) {
                        mbar_wait(bar_empty + stage * 8, phase ^ 1);
                        const uint32_t full = bar_full + stage * 8;
                        mbar_arrive_expect_tx(full, STAGE_BYTES);
                        const uint32_t s0 = sbase + smem_layout::OPERANDS + stage * STAGE_BYTES;
                        if (half == 0) {
                            tma_load_2d(s0, &map_hi, full, kc, row_a, keep);                          // A_hi  128 rows
                            tma_load_2d(s0 + A_BYTES, &map_hi, full, kc, row_b, keep);                // B_hi  rows 0..127
                            tma_load_2d(s0 + 2 * A_BYTES, &map_hi, full, kc, row_b + 128, keep);      // B_hi  rows 128..255
                        } else if constexpr (FP8X) {
                            const int kc8 = kb * 128;  // bytes: [hi8 | lo8] of this K block
                            tma_load_2d(s0, &map_lo, full, kc8, row_a, keep);                         // A_8
                            tma_load_2d(s0 + A_BYTES, &map_b, full, kc8, row_b, keep);                // B_8
                            tma_load_2d(s0 + 2 * A_BYTES, &map_b, full, kc8, row_b + 128, keep);
                        } else {
                            tma_load_2d(s0, &map_lo, full, kc, row_a, keep);                          // A_lo
                            tma_load_2d(s0 + A_BYTES, &map_lo, full, kc, row_b, keep);                // B_lo
                            tma_load_2d(s0 + 2 * A_BYTES, &map_lo, full, kc, row_b + 128, keep);
                        }
                        if (++stage == STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ============================== MMA issuer ==============================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_f16(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int t = first_tile; t < p.num_tiles; t += tile_step, ++it) {
                const int buf = it & 1;
                const uint32_t use = static_cast<uint32_t>(it >> 1);
                mbar_wait(bar_tempty + buf * 8, (use & 1) ^ 1);  // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + buf * BN;
                for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                    // slot 0: (A_hi, B_hi)
                    const int st0 = stage;
                    mbar_wait(bar_full + st0 * 8, phase);
                    tc_fence_after();
                    const uint32_t s0 = sbase + smem_layout::OPERANDS + st0 * STAGE_BYTES;
                    const uint64_t a_hi = make_desc_sw128(s0), b_hi = make_desc_sw128(s0 + A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t adv = static_cast<uint64_t>((k * UMMA_K * 2) >> 4);  // +32 bytes per K step inside the swizzle row
                        tc_mma_f16(tmem_d, a_hi + adv, b_hi + adv, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    if constexpr (FP8X) { tc_commit(bar_empty + st0 * 8); }  // frees the slot when these MMAs retire
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                    // slot 1: cross-term operands
                    const int st1 = stage;
                    mbar_wait(bar_full + st1 * 8, phase);
                    tc_fence_after();
                    const uint32_t s1 = sbase + smem_layout::OPERANDS + st1 * STAGE_BYTES;
                    const uint64_t a_lo = make_desc_sw128(s1), b_lo = make_desc_sw128(s1 + A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t adv = static_cast<uint64_t>((k * UMMA_K * 2) >> 4);
                        if constexpr (FP8X) {
                            // same 32-byte step: 32 e4m3 elements of [hi8 | lo8] x [lo8 | hi8]^T (idesc formats 0 = E4M3, same bit layout)
                            tc_mma_f8(tmem_d, a_lo + adv, b_lo + adv, idesc, 1u);
                        } else {
                            tc_mma_f16(tmem_d, a_hi + adv, b_lo + adv, idesc, 1u);
                            tc_mma_f16(tmem_d, a_lo + adv, b_hi + adv, idesc, 1u);
                        }
                    }
                    if constexpr (!FP8X) { tc_commit(bar_empty + st0 * 8); }
                    tc_commit(bar_empty + st1 * 8);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                tc_commit(bar_tfull + buf * 8);  // accumulator complete -> epilogue
            }
        }
    } else {
        // ============================== epilogue warps (2..5) ==============================
        const int q = warp & 3;                 // TMEM lane quadrant this warp may access
        const int half = (warp - 2) >> 2;       // column half of the 256-wide tile this warp evaluates
        const int et = threadIdx.x - 64;        // 0..255
        float *rowred = reinterpret_cast<float *>(smem + smem_layout::ROWRED);
        float4 *colbuf = reinterpret_cast<float4 *>(smem + smem_layout::COLBUF);
        float *colred = reinterpret_cast<float *>(smem + smem_layout::COLRED);
        int it = 0;
        for (int t = first_tile; t < p.num_tiles; t += tile_step, ++it) {
            const int buf = it & 1;
            const uint32_t use = static_cast<uint32_t>(it >> 1);
            const int2 tile = p.tiles[t];
            const int64_t row0 = static_cast<int64_t>(tile.x) * BM, col0 = static_cast<int64_t>(tile.y) * BN;
            // stage the per-column data of this tile
            colbuf[buf * BN + et] = __ldg(&p.pk[col0 + et]);
            const int r = q * 32 + lane;
            const int64_t gi = row0 + r;
            const float4 mine = __ldg(&p.pk[gi]);  // {c_i, q_i, d_i, -}
            const float ai = p.qa - mine.y;
            const float di = mine.z;
            const bool diag_tile = (row0 < col0 + BN);  // tile touches or crosses the diagonal

            mbar_wait(bar_tfull + buf * 8, use & 1);
            tc_fence_after();
            asm volatile("bar.sync 1, 256;" ::: "memory");  // colbuf visible to the epilogue warps

            float rowsum = 0.f;
            float *cr = colred + (buf * 4 + q) * BN;
#pragma unroll 1
            for (int g = half * (BN / 64); g < (half + 1) * (BN / 64); ++g) {
                uint32_t u[32];
                tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(buf * BN + g * 32), u);
                tmem_ld_wait();
                float v[32];
                if (!diag_tile) {
#pragma unroll
                    for (int cc = 0; cc < 32; ++cc) {
                        const float4 cj = colbuf[buf * BN + g * 32 + cc];
                        const float kv = kernel_from_gram<KERNEL>(__uint_as_float(u[cc]), mine.x, cj.x, p.w, p.coef0, p.degree);
                        if constexpr (STORE) { u[cc] = __float_as_uint(kv); }
                        const float tv = kv + (ai - cj.y);
                        rowsum = fmaf(tv, cj.z, rowsum);
                        v[cc] = __fmul_rn(tv, di);  // a separate multiply in every instantiation: the STORE pass returns the recompute pass's bits
                    }
                } else {
#pragma unroll
                    for (int cc = 0; cc < 32; ++cc) {
                        const float4 cj = colbuf[buf * BN + g * 32 + cc];
                        const int64_t gj = col0 + g * 32 + cc;
                        const float kv = kernel_from_gram<KERNEL>(__uint_as_float(u[cc]), mine.x, cj.x, p.w, p.coef0, p.degree);
                        if constexpr (STORE) { u[cc] = __float_as_uint(kv); }   // the stored tile is complete (no triangle mask)
                        const float tv = kv + (ai - cj.y);
                        rowsum = gi >= gj ? fmaf(tv, cj.z, rowsum) : rowsum;  // lower triangle incl. diagonal -> row i
                        v[cc] = gi > gj ? __fmul_rn(tv, di) : 0.f;                        // strictly lower -> mirrored into column j
                    }
                }
                if constexpr (STORE) {
                    // the 32 lanes of the warp hold 32 consecutive rows: every store instruction writes 512 contiguous bytes (kc_quad_index)
                    float4 *dst = reinterpret_cast<float4 *>(p.Kc) + static_cast<int64_t>(t) * (BM * BN / 4) + kc_quad_index(g * 8, r);
#pragma unroll
                    for (int c4 = 0; c4 < 8; ++c4) {
                        __stcs(dst + c4 * BM, make_float4(__uint_as_float(u[c4 * 4]), __uint_as_float(u[c4 * 4 + 1]), __uint_as_float(u[c4 * 4 + 2]),
                                                          __uint_as_float(u[c4 * 4 + 3])));
                    }
                }
                const float cs = warp_column_sums(v, lane);
                cr[g * 32 + lane] = cs;
            }
            // accumulator fully read: hand the TMEM buffer back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { mbar_arrive(bar_tempty + buf * 8); }
            if (half == 1) { rowred[buf * BM + r] = rowsum; }
            asm volatile("bar.sync 1, 256;" ::: "memory");  // colred / rowred complete
            if (half == 0) { p.Pr[static_cast<int64_t>(tile.y) * p.ldP + gi] = rowsum + rowred[buf * BM + r]; }
            {
                const float *c0 = colred + buf * 4 * BN;
                p.Pc[static_cast<int64_t>(tile.x) * p.ldP + col0 + et] = (c0[et] + c0[BN + et]) + (c0[2 * BN + et] + c0[3 * BN + et]);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Cached kernel matrix: the matvec from the stored tiles — an HBM-bound symmetric GEMV.
//   CENTRE:   t_ij = K_ij + QA_cost - q_i - q_j   (dense data: the reference's per-entry centring, svm_kernel.cpp:32)
//   !CENTRE:  t_ij = S_ij                          (sparse rows with a densely filled S: the rank-1 part is added in the O(m) epilogue)
//   row part  Pr[J][i]    = sum_j t_ij w_j  (i >= j)   one partial vector per column strip J
//   col part  Pc[run][j]  = sum_i t_ij w_i  (i > j)    one partial vector per RUN of tiles of a strip
// The rank's tiles are sorted by (strip J, row block I) and cut into equal chunks, one per CTA: a CTA walks down a strip and keeps
// the 32 mirrored column sums of every thread in registers across the tiles of the run — ONE shuffle butterfly per run instead of
// one per tile — so a strip of up to 512 tiles leaves a handful of column partial vectors (runs per strip <= strip length /
// chunk length + 2) instead of one per tile; the reduction kernel reads ~100x fewer column partials than before.
// Tile layout = kc_quad_index (quad-column-major): warp w owns the 32 columns [32 w, 32 w + 32) over all 128 rows, lane l the rows
// l, 32 + l, 64 + l, 96 + l, so every load instruction of a warp streams 512 contiguous bytes and the warp's 16 KB are one
// contiguous range.  Row partials: four registers per thread, combined over the eight warps through (double-buffered) shared
// memory in a fixed tree.  Fixed order everywhere => bit-reproducible.
// ---------------------------------------------------------------------------------------------------------------------
struct sweep_params {
    const float *Kc;         // stored tiles, [storage index][BN / 4][BM] float4
    const int4 *list;        // tiles of this rank sorted by (J, I): {I, J, storage index, 0}
    int num_tiles;
    int nchunks;
    const int4 *chunks;      // [nchunks] {first tile, end tile, partial-vector slot of the chunk's first run, first slot of the chunk's group}
    const float4 *pk;        // [n_pad] {-, q_i, w_i, -}
    float *Pr;               // [T_c][ldP]
    float *Pc;               // [slots][ldP]: one per (group, run of a strip)
    int64_t ldP;
    float qa;
    const float *done;
};

template <bool CENTRE>
__global__ void __launch_bounds__(256, 2) k_tile_sweep(const sweep_params p) {
    if (p.done != nullptr && *p.done != 0.f) { return; }
    __shared__ float4 colq[BN / 4];      // q_j of the strip's columns, one float4 per column quad
    __shared__ float4 cold[BN / 4];      // w_j
    __shared__ float rowpart[2][8][BM];  // per warp: totals of the 128 rows over its 32 columns (double-buffered over the tiles)
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    for (int c = blockIdx.x; c < p.nchunks; c += gridDim.x) {
        const int4 chunk = __ldg(&p.chunks[c]);
        const int k0 = chunk.x, k1 = chunk.y;
        int curJ = -1, slot = 0;
        float cs[32];
        for (int k = k0; k < k1; ++k) {
            const int4 tile = __ldg(&p.list[k]);
            const int64_t row0 = static_cast<int64_t>(tile.x) * BM, col0 = static_cast<int64_t>(tile.y) * BN;
            if (tile.y != curJ) {  // block-uniform: a new strip (the tile barrier below has retired every reader of colq / cold)
                if (curJ >= 0) {
                    const float ctot = warp_column_sums(cs, lane);  // lane j: column 32 w + j over the rows of the run
                    p.Pc[static_cast<int64_t>(slot) * p.ldP + static_cast<int64_t>(curJ) * BN + 32 * w + lane] = ctot;
                }
                slot = k == k0 ? chunk.z : chunk.w;  // (a run that starts inside a chunk starts its strip within the group)
                curJ = tile.y;
#pragma unroll
                for (int i = 0; i < 32; ++i) { cs[i] = 0.f; }
                const float4 v = __ldg(&p.pk[col0 + t]);
                reinterpret_cast<float *>(colq)[t] = v.y;
                reinterpret_cast<float *>(cold)[t] = v.z;
                __syncthreads();
            }
            const bool diag_tile = row0 < col0 + BN;
            float ai[4], di[4];  // (QA_cost - q_i, w_i) of this thread's four rows
#pragma unroll
            for (int rg = 0; rg < 4; ++rg) {
                const float4 v = __ldg(&p.pk[row0 + rg * 32 + lane]);
                ai[rg] = p.qa - v.y;
                di[rg] = v.z;
            }
            const float4 *base = reinterpret_cast<const float4 *>(p.Kc) + static_cast<int64_t>(tile.z) * (BM * BN / 4) + kc_quad_index(8 * w, lane);
            float rs[4] = { 0.f, 0.f, 0.f, 0.f };
#pragma unroll
            for (int q0 = 0; q0 < 8; q0 += 2) {
                float4 kv[2][4];
#pragma unroll
                for (int kk = 0; kk < 2; ++kk) {
#pragma unroll
                    for (int rg = 0; rg < 4; ++rg) { kv[kk][rg] = __ldcs(base + (q0 + kk) * BM + rg * 32); }  // eight 16-byte loads in flight
                }
#pragma unroll
                for (int kk = 0; kk < 2; ++kk) {
                    const int cq = 8 * w + q0 + kk;
                    const float4 q4 = colq[cq], d4 = cold[cq];  // same address for the whole warp: broadcast
                    const float qj[4] = { q4.x, q4.y, q4.z, q4.w };
                    const float dj[4] = { d4.x, d4.y, d4.z, d4.w };
                    float ca[4] = { 0.f, 0.f, 0.f, 0.f };
#pragma unroll
                    for (int rg = 0; rg < 4; ++rg) {
                        const float e[4] = { kv[kk][rg].x, kv[kk][rg].y, kv[kk][rg].z, kv[kk][rg].w };
                        if (!diag_tile) {
#pragma unroll
                            for (int cc = 0; cc < 4; ++cc) {
                                const float tv = CENTRE ? e[cc] + (ai[rg] - qj[cc]) : e[cc];
                                rs[rg] = fmaf(tv, dj[cc], rs[rg]);
                                ca[cc] = fmaf(tv, di[rg], ca[cc]);
                            }
                        } else {
                            const int64_t gi = row0 + rg * 32 + lane;
#pragma unroll
                            for (int cc = 0; cc < 4; ++cc) {
                                const int64_t gj = col0 + 4 * cq + cc;
                                const float tv = CENTRE ? e[cc] + (ai[rg] - qj[cc]) : e[cc];
                                rs[rg] = gi >= gj ? fmaf(tv, dj[cc], rs[rg]) : rs[rg];  // lower triangle incl. diagonal -> row i
                                ca[cc] = gi > gj ? fmaf(tv, di[rg], ca[cc]) : ca[cc];   // strictly lower -> mirrored into column j
                            }
                        }
                    }
#pragma unroll
                    for (int cc = 0; cc < 4; ++cc) { cs[(q0 + kk) * 4 + cc] += ca[cc]; }
                }
            }
            const int buf = (k - k0) & 1;
#pragma unroll
            for (int rg = 0; rg < 4; ++rg) { rowpart[buf][w][rg * 32 + lane] = rs[rg]; }
            __syncthreads();  // one barrier per tile: rowpart[buf] complete; nobody still reads this tile's colq / cold
            if (t < BM) {
                const float(*rp)[BM] = rowpart[buf];
                p.Pr[static_cast<int64_t>(tile.y) * p.ldP + row0 + t] =
                    ((rp[0][t] + rp[1][t]) + (rp[2][t] + rp[3][t])) + ((rp[4][t] + rp[5][t]) + (rp[6][t] + rp[7][t]));
            }
        }
        if (curJ >= 0) {
            const float ctot = warp_column_sums(cs, lane);
            p.Pc[static_cast<int64_t>(slot) * p.ldP + static_cast<int64_t>(curJ) * BN + 32 * w + lane] = ctot;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// helper kernels: max |x|, fp16 hi/lo split, per-index pack, partial-vector reduction
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_absmax(const float *__restrict__ X, int64_t count, unsigned int *__restrict__ out) {
    float mx = 0.f;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t) gridDim.x * blockDim.x) {
        mx = fmaxf(mx, fabsf(X[i]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o)); }
    if ((threadIdx.x & 31) == 0) { atomicMax(out, __float_as_uint(mx)); }  // non-negative floats order like their bit patterns
}

// hi = fp16(x s), lo = fp16(x s - hi); rows >= n and columns >= ldx are zero.
// FP8X planes (when pa != nullptr): per 64-element K block 128 bytes, pa = [e4m3(hi / 64) | e4m3(lo * 64)], pb = [e4m3(lo * 64) | e4m3(hi / 64)]
// (hi < 2^14 -> hi / 64 < 256; |lo| <= 4 -> lo * 64 <= 256; e4m3 holds +-448).
__global__ void __launch_bounds__(256) k_split_fp16(const float *__restrict__ X, int64_t n, int64_t ldx, float s, __half *__restrict__ hi,
                                                    __half *__restrict__ lo, uint8_t *__restrict__ pa, uint8_t *__restrict__ pb, int64_t n_pad,
                                                    int64_t ldh) {
    const int64_t total = n_pad * ldh;
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t) gridDim.x * blockDim.x) {
        const int64_t row = e / ldh, col = e - row * ldh;
        float x = 0.f;
        if (row < n && col < ldx) { x = X[row * ldx + col] * s; }
        const __half h = __float2half_rn(x);
        const float l = x - __half2float(h);
        hi[e] = h;
        if (lo != nullptr) { lo[e] = __float2half_rn(l); }
        if (pa != nullptr) {
            const uint8_t h8 = (uint8_t) __nv_cvt_float_to_fp8(__half2float(h) * 0.015625f, __NV_SATFINITE, __NV_E4M3);
            const uint8_t l8 = (uint8_t) __nv_cvt_float_to_fp8(l * 64.f, __NV_SATFINITE, __NV_E4M3);
            const int64_t base = row * (2 * ldh) + (col >> 6) * 128 + (col & 63);
            pa[base] = h8;
            pa[base + 64] = l8;
            pb[base] = l8;
            pb[base + 64] = h8;
        }
    }
}

// pk[i] = {c_i, q_i, d_i, 0} for i < m, zeros above; c_i = -gamma log2(e) |x_i|^2 (rbf) or 0
__global__ void __launch_bounds__(256) k_pack(const float *__restrict__ norms, const float *__restrict__ q, const float *__restrict__ d,
                                              int64_t m, int64_t n_pad, float cscale, float4 *__restrict__ pk) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pad) { return; }
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < m) { v = make_float4(cscale * norms[i], q[i], d[i], 0.f); }
    pk[i] = v;
}

// ---------------------------------------------------------------------------------------------------------------------
// The same sweep for DOUBLE tiles (dense double polynomial / rbf): 128 x 256 tiles of 8-byte kernel values in
// pair-column-major order (double2 of columns (2 cp, 2 cp + 1), row r at cp * 128 + r), 512 threads: warp w owns the 16 columns
// [16 w, 16 w + 16), lane l the rows l, 32 + l, 64 + l, 96 + l — every load instruction streams 512 contiguous bytes.  Same chunk /
// run / partial-vector structure as k_tile_sweep.  t_ij = K_ij + QA_cost - q_i - q_j.
// ---------------------------------------------------------------------------------------------------------------------
struct sweep_params_f64 {
    const double *Kc;
    const int4 *list;
    int num_tiles;
    int nchunks;
    const int4 *chunks;
    const double2 *pk;       // [n_pad] {q_i, w_i}
    double *Pr;
    double *Pc;
    int64_t ldP;
    double qa;
    const double *done;
};

__device__ __forceinline__ double warp_column_sums16(double (&v)[16], int lane) {
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) {
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int k = 0; k < o; ++k) {
            const double send = up ? v[k] : v[k + o];
            const double keep = up ? v[k + o] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);  // both half-warps end with column (lane & 15)
}

__global__ void __launch_bounds__(512, 1) k_tile_sweep_f64(const sweep_params_f64 p) {
    if (p.done != nullptr && *p.done != 0.0) { return; }
    __shared__ double colq[BN], cold[BN];
    __shared__ double rowpart[2][16][BM];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    for (int c = blockIdx.x; c < p.nchunks; c += gridDim.x) {
        const int4 chunk = __ldg(&p.chunks[c]);
        const int k0 = chunk.x, k1 = chunk.y;
        int curJ = -1, slot = 0;
        double cs[16];
        for (int k = k0; k < k1; ++k) {
            const int4 tile = __ldg(&p.list[k]);
            const int64_t row0 = static_cast<int64_t>(tile.x) * BM, col0 = static_cast<int64_t>(tile.y) * BN;
            if (tile.y != curJ) {
                if (curJ >= 0) {
                    const double ctot = warp_column_sums16(cs, lane);
                    if (lane < 16) { p.Pc[static_cast<int64_t>(slot) * p.ldP + static_cast<int64_t>(curJ) * BN + 16 * w + lane] = ctot; }
                }
                slot = k == k0 ? chunk.z : chunk.w;  // (a run that starts inside a chunk starts its strip within the group)
                curJ = tile.y;
#pragma unroll
                for (int i = 0; i < 16; ++i) { cs[i] = 0.0; }
                if (t < BN) {
                    const double2 v = __ldg(&p.pk[col0 + t]);
                    colq[t] = v.x;
                    cold[t] = v.y;
                }
                __syncthreads();
            }
            const bool diag_tile = row0 < col0 + BN;
            double ai[4], di[4];
#pragma unroll
            for (int rg = 0; rg < 4; ++rg) {
                const double2 v = __ldg(&p.pk[row0 + rg * 32 + lane]);
                ai[rg] = p.qa - v.x;
                di[rg] = v.y;
            }
            const double2 *base = reinterpret_cast<const double2 *>(p.Kc) + static_cast<int64_t>(tile.z) * (BM * BN / 2) + static_cast<int64_t>(8 * w) * BM + lane;
            double rs[4] = { 0.0, 0.0, 0.0, 0.0 };
#pragma unroll
            for (int q0 = 0; q0 < 8; q0 += 2) {
                double2 kv[2][4];
#pragma unroll
                for (int kk = 0; kk < 2; ++kk) {
#pragma unroll
                    for (int rg = 0; rg < 4; ++rg) { kv[kk][rg] = __ldcs(base + (q0 + kk) * BM + rg * 32); }
                }
#pragma unroll
                for (int kk = 0; kk < 2; ++kk) {
                    const int cp = 8 * w + q0 + kk;  // column pair inside the tile
                    const double qj[2] = { colq[2 * cp], colq[2 * cp + 1] };
                    const double dj[2] = { cold[2 * cp], cold[2 * cp + 1] };
                    double ca[2] = { 0.0, 0.0 };
                    if (!diag_tile) {
#pragma unroll
                        for (int rg = 0; rg < 4; ++rg) {
                            const double e[2] = { kv[kk][rg].x, kv[kk][rg].y };
#pragma unroll
                            for (int cc = 0; cc < 2; ++cc) {
                                const double tv = e[cc] + (ai[rg] - qj[cc]);
                                rs[rg] = fma(tv, dj[cc], rs[rg]);
                                ca[cc] = fma(tv, di[rg], ca[cc]);
                            }
                        }
                    } else {
                        const int off = (int) (col0 - row0);  // row i of the tile is below column j  <=>  i - j > off
#pragma unroll
                        for (int rg = 0; rg < 4; ++rg) {
                            const double e[2] = { kv[kk][rg].x, kv[kk][rg].y };
#pragma unroll
                            for (int cc = 0; cc < 2; ++cc) {
                                const int dif = rg * 32 + lane - (2 * cp + cc);
                                const double tv = e[cc] + (ai[rg] - qj[cc]);
                                rs[rg] = dif >= off ? fma(tv, dj[cc], rs[rg]) : rs[rg];  // lower triangle incl. diagonal -> row i
                                ca[cc] = dif > off ? fma(tv, di[rg], ca[cc]) : ca[cc];   // strictly lower -> mirrored into column j
                            }
                        }
                    }
                    cs[(q0 + kk) * 2] += ca[0];
                    cs[(q0 + kk) * 2 + 1] += ca[1];
                }
            }
            const int buf = (k - k0) & 1;
#pragma unroll
            for (int rg = 0; rg < 4; ++rg) { rowpart[buf][w][rg * 32 + lane] = rs[rg]; }
            __syncthreads();
            if (t < BM) {
                const double(*rp)[BM] = rowpart[buf];
                double s = 0.0;
#pragma unroll
                for (int ww = 0; ww < 16; ww += 4) { s += (rp[ww][t] + rp[ww + 1][t]) + (rp[ww + 2][t] + rp[ww + 3][t]); }
                p.Pr[static_cast<int64_t>(tile.y) * p.ldP + row0 + t] = s;
            }
        }
        if (curJ >= 0) {
            const double ctot = warp_column_sums16(cs, lane);
            if (lane < 16) { p.Pc[static_cast<int64_t>(slot) * p.ldP + static_cast<int64_t>(curJ) * BN + 16 * w + lane] = ctot; }
        }
    }
}

// pk[i] = {q_i, w_i} for i < m, zeros above
__global__ void __launch_bounds__(256) k_pack_f64(const double *__restrict__ q, const double *__restrict__ w, int64_t m, int64_t n_pad, double2 *__restrict__ pk) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pad) { return; }
    pk[i] = i < m ? make_double2(q[i], w[i]) : make_double2(0.0, 0.0);
}

// Sum of the partial vectors of element i -> out[i], or (several ranks, peer exchange) straight into slot `rank` of every
// peer's exchange buffer over NVLink: the reduction kernel IS the producer of the exchange (xchg.cuh).
template <typename T>
struct reduce_dst_t {
    T *out;
    xchg_push push;
    int remote;  // 0: out, 1: peers
};
using reduce_dst = reduce_dst_t<float>;
template <typename T>
__device__ __forceinline__ void reduce_emit(const reduce_dst_t<T> &dst, unsigned long long seq, int64_t i, T v) {
    if (dst.remote) {
#pragma unroll 1
        for (int r = 0; r < dst.push.world; ++r) { xchg_slot<T>(dst.push, r, seq, dst.push.rank)[i] = v; }
    } else {
        dst.out[i] = v;
    }
}

// CTAs of a reduction kernel: 32 * 8 / groups elements each (see k_sweep_reduce)
inline unsigned reduce_grid(int64_t m, int groups) {
    const int64_t per = 32 * (8 / groups);
    return (unsigned) ((m + per - 1) / per);
}

// Fixed summation tree over the partial sums of up to eight tile GROUPS (DESIGN.md §6 "N-independent results"): the tiles are dealt
// to eight canonical groups; a rank of a world of 1 / 2 / 4 / 8 owns 8 / 4 / 2 / 1 consecutive groups, sums each group's partial
// vectors in the group's own fixed order, combines its groups by the balanced tree below, and the consumer combines the ranks' slots
// by the same tree — the value of every element is bit-identical for 1, 2, 4 and 8 GPUs.
template <typename T>
__device__ __forceinline__ T group_tree(const T (&p)[8], int n) {
    switch (n) {
        case 8: return ((p[0] + p[1]) + (p[2] + p[3])) + ((p[4] + p[5]) + (p[6] + p[7]));
        case 4: return (p[0] + p[1]) + (p[2] + p[3]);
        case 2: return p[0] + p[1];
        default: {
            T s = p[0];
            for (int g = 1; g < n; ++g) { s += p[g]; }
            return s;
        }
    }
}

// Partial vectors of the Gram kernel (one per tile): out[i] = sum over the tiles (I_i, J) of Pr[J][i] + sum over the tiles (I, J_i) of
// Pc[I][i], per group (gmap[I][J] = local group of the tile, 255 = not this rank's) in ascending J, then ascending I; one warp per
// group, 32 elements per CTA.
template <typename T>
__global__ void __launch_bounds__(256) k_tc_reduce(const T *__restrict__ Pr, const T *__restrict__ Pc, int64_t ldP, int64_t m, int T_r, int T_c,
                                                   const uint8_t *__restrict__ gmap, int groups, const reduce_dst_t<T> dst, const T *__restrict__ done) {
    __shared__ bool is_last;
    __shared__ T part[8][32];
    if (done != nullptr && *done != T(0)) { return; }
    const unsigned long long seq = dst.remote ? xchg_next_seq(dst.push) : 0ull;
    // groups in {1, 2, 4, 8}: the eight warps serve 8 / groups sets of 32 consecutive elements, one warp per (set, group)
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, w = wid % groups, set = wid / groups;
    const int64_t i = ((int64_t) blockIdx.x * (8 / groups) + set) * 32 + lane;
    T tot = T(0);
    if (i < m) {
        const int Ii = (int) (i / BM), Ji = (int) (i / BN);  // (the same for the 32 elements of the warp: the map reads are warp-uniform)
        T s0 = T(0), s1 = T(0);
        for (int J = 0; J <= Ii / 2; ++J) {
            if (gmap[Ii * T_c + J] == w) { s0 += Pr[(int64_t) J * ldP + i]; }
        }
        for (int I = 2 * Ji; I < T_r; ++I) {
            if (gmap[I * T_c + Ji] == w) { s1 += Pc[(int64_t) I * ldP + i]; }
        }
        tot = s0 + s1;
    }
    part[wid][lane] = tot;
    __syncthreads();
    if (w == 0 && i < m) {
        T pg[8];
#pragma unroll
        for (int g = 0; g < 8; ++g) { pg[g] = g < groups ? part[set * groups + g][lane] : T(0); }
        reduce_emit(dst, seq, i, group_tree(pg, groups));
    }
    if (dst.remote) { xchg_publish(dst.push, seq, &is_last); }
}

// Partial vectors of the sweep: per group g, the row partials of the group's column strips in row block I_i
// (jlist[jptr[I (G + 1) + g] .. jptr[I (G + 1) + g + 1])) with four interleaved accumulators, then the column partials of the group's
// runs of strip J_i (slots slot_base[g] .. + nruns[g T_c + J]).  One warp per group, 32 elements per CTA, groups combined by group_tree.
template <typename T>
__global__ void __launch_bounds__(256) k_sweep_reduce(const T *__restrict__ Pr, const T *__restrict__ Pc, int64_t ldP, int64_t m, int T_c,
                                                      const int *__restrict__ jptr, const int *__restrict__ jlist, const int *__restrict__ nruns,
                                                      const int *__restrict__ slot_base, int groups, const reduce_dst_t<T> dst, const T *__restrict__ done) {
    __shared__ bool is_last;
    __shared__ T part[8][32];
    if (done != nullptr && *done != T(0)) { return; }
    const unsigned long long seq = dst.remote ? xchg_next_seq(dst.push) : 0ull;
    // groups in {1, 2, 4, 8}: the eight warps serve 8 / groups sets of 32 consecutive elements, one warp per (set, group)
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, w = wid % groups, set = wid / groups;
    const int64_t i = ((int64_t) blockIdx.x * (8 / groups) + set) * 32 + lane;
    T tot = T(0);
    if (i < m) {
        const int Ii = (int) (i / BM), Ji = (int) (i / BN);  // (the same for the 32 elements of the warp)
        const int ka = __ldg(&jptr[Ii * (groups + 1) + w]), kb = __ldg(&jptr[Ii * (groups + 1) + w + 1]);
        T s0 = T(0), s1 = T(0), s2 = T(0), s3 = T(0);
        int k = ka;
        for (; k + 4 <= kb; k += 4) {
            const int Ja = __ldg(&jlist[k]), Jb = __ldg(&jlist[k + 1]), Jc = __ldg(&jlist[k + 2]), Jd = __ldg(&jlist[k + 3]);
            const T a0 = Pr[(int64_t) Ja * ldP + i], a1 = Pr[(int64_t) Jb * ldP + i], a2 = Pr[(int64_t) Jc * ldP + i], a3 = Pr[(int64_t) Jd * ldP + i];
            s0 += a0;
            s1 += a1;
            s2 += a2;
            s3 += a3;
        }
        for (; k < kb; ++k) { s0 += Pr[(int64_t) __ldg(&jlist[k]) * ldP + i]; }
        T c = T(0);
        const int nr = __ldg(&nruns[w * T_c + Ji]), sb = __ldg(&slot_base[w]);
        for (int r = 0; r < nr; ++r) { c += Pc[(int64_t) (sb + r) * ldP + i]; }
        tot = ((s0 + s1) + (s2 + s3)) + c;
    }
    part[wid][lane] = tot;
    __syncthreads();
    if (w == 0 && i < m) {
        T pg[8];
#pragma unroll
        for (int g = 0; g < 8; ++g) { pg[g] = g < groups ? part[set * groups + g][lane] : T(0); }
        reduce_emit(dst, seq, i, group_tree(pg, groups));
    }
    if (dst.remote) { xchg_publish(dst.push, seq, &is_last); }
}

}  // namespace tc

// ---------------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------------
// sweep over stored tiles: the rank's tiles sorted by (strip, row block), cut into chunks (one per CTA), and what the reduction
// kernel needs to find the partial vectors (k_tile_sweep / k_sweep_reduce)
struct sweep_plan {
    dbuf list, chunks, nruns, slot_base, jptr, jlist;
    int nchunks{ 0 };
    int groups{ 1 };
    int slots{ 1 };  // column partial vectors (rows of Pc) in use
};

// canonical tile groups (k_sweep_reduce / group_tree): 8 when the world divides 8, else one per rank (no cross-world reproducibility)
int canonical_groups(int world) { return world >= 1 && 8 % world == 0 ? 8 : world; }

// tiles: this rank's tiles in storage order; grp[k] = local group of tile k (non-decreasing, 0 .. groups - 1)
void build_sweep_plan(ctx &c, sweep_plan &sw, const std::vector<int2> &tiles, const std::vector<int> &grp, int groups, int T_r, int T_c) {
    const int nt = (int) tiles.size();
    sw.groups = groups;
    std::vector<int4> list((size_t) nt);
    for (int t = 0; t < nt; ++t) { list[(size_t) t] = make_int4(tiles[(size_t) t].x, tiles[(size_t) t].y, t, grp[(size_t) t]); }
    std::sort(list.begin(), list.end(), [](const int4 &a, const int4 &b) { return a.w != b.w ? a.w < b.w : (a.y != b.y ? a.y < b.y : a.x < b.x); });
    // chunks: every group is cut into the same number of equal chunks whatever the world size is (one CTA per chunk; 296 = one wave
    // of two CTAs per SM for a rank that owns a single group)
    static const int CHUNKS_PER_GROUP = [] {
        const char *e = std::getenv("B200SVM_SWEEP_CHUNKS");  // experiment switch; results are comparable across world sizes only for equal values
        const int v = e != nullptr ? std::atoi(e) : 0;
        return v > 0 ? v : 2 * NUM_SMS;
    }();
    std::vector<int4> chunks;
    std::vector<int> nruns((size_t) groups * (size_t) T_c, 0), slot_base((size_t) groups, 0);
    int g0 = 0, slots = 0;
    for (int g = 0; g < groups; ++g) {
        int g1 = g0;
        while (g1 < nt && list[(size_t) g1].w == g) { ++g1; }
        const int ng = g1 - g0, nch = std::max(1, std::min(ng, CHUNKS_PER_GROUP));
        slot_base[(size_t) g] = slots;
        int max_runs = 0;
        for (int ch = 0; ch < nch && ng > 0; ++ch) {
            const int k0 = g0 + (int) ((int64_t) ng * ch / nch), k1 = g0 + (int) ((int64_t) ng * (ch + 1) / nch);
            int curJ = -1, first = 0;
            for (int k = k0; k < k1; ++k) {
                const int J = list[(size_t) k].y;
                if (J != curJ) {
                    int &cnt = nruns[(size_t) g * (size_t) T_c + (size_t) J];
                    if (k == k0) { first = cnt; }
                    ++cnt;
                    max_runs = std::max(max_runs, cnt);
                    curJ = J;
                }
            }
            chunks.push_back(make_int4(k0, k1, slots + first, slots));
        }
        slots += std::max(max_runs, 1);
        g0 = g1;
    }
    sw.nchunks = (int) chunks.size();
    sw.slots = std::max(slots, 1);
    for (int4 &e : list) { e.w = 0; }
    // column strips per (row block, group), ascending
    std::vector<std::vector<int>> per((size_t) T_r * (size_t) groups);
    {
        std::vector<int4> bygroup((size_t) nt);
        for (int t = 0; t < nt; ++t) { bygroup[(size_t) t] = make_int4(tiles[(size_t) t].x, tiles[(size_t) t].y, grp[(size_t) t], 0); }
        std::sort(bygroup.begin(), bygroup.end(), [](const int4 &a, const int4 &b) { return a.y < b.y; });
        for (const int4 &e : bygroup) { per[(size_t) e.x * (size_t) groups + (size_t) e.z].push_back(e.y); }
    }
    std::vector<int> jptr((size_t) T_r * (size_t) (groups + 1), 0), jlist;
    jlist.reserve((size_t) nt);
    for (int I = 0; I < T_r; ++I) {
        for (int g = 0; g < groups; ++g) {
            jptr[(size_t) I * (size_t) (groups + 1) + (size_t) g] = (int) jlist.size();
            const std::vector<int> &v = per[(size_t) I * (size_t) groups + (size_t) g];
            jlist.insert(jlist.end(), v.begin(), v.end());
        }
        jptr[(size_t) I * (size_t) (groups + 1) + (size_t) groups] = (int) jlist.size();
    }
    const auto up = [&](dbuf &dst, const void *src, size_t bytes) {
        dst.reserve(std::max<size_t>(bytes, 16));
        if (bytes != 0) { B200_CUDA_CHECK(cudaMemcpyAsync(dst.p, src, bytes, cudaMemcpyHostToDevice, c.stream)); }
    };
    up(sw.list, list.data(), list.size() * sizeof(int4));
    up(sw.chunks, chunks.data(), chunks.size() * sizeof(int4));
    up(sw.nruns, nruns.data(), nruns.size() * sizeof(int));
    up(sw.slot_base, slot_base.data(), slot_base.size() * sizeof(int));
    up(sw.jptr, jptr.data(), jptr.size() * sizeof(int));
    up(sw.jlist, jlist.data(), jlist.size() * sizeof(int));
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));  // the host vectors must outlive the copies
}

namespace {

struct tc_state {
    CUtensorMap map_hi{}, map_lo{}, map_a8{}, map_b8{};
    dbuf tiles, pk, Pr, Pc, amax, P8a, P8b, own, Kc;
    sweep_plan sweep;
    int groups{ 1 };  // canonical tile groups of this rank (group_tree)
    bool kc_valid{ false }, kc_refused{ false };  // cached kernel-matrix tiles of this rank
    bool fp8x{ false };
    int num_tiles{ 0 };
    int T_r{ 0 }, T_c{ 0 };
    int64_t ldP{ 0 };
    bool ready{ false };
};

// one tc_state per handle, owned through ctx::tc_opaque
tc_state &state_of(ctx &c) {
    if (c.tc_opaque == nullptr) {
        c.tc_opaque = new tc_state();
        c.tc_opaque_free = [](void *p) { delete static_cast<tc_state *>(p); };
    }
    return *static_cast<tc_state *>(c.tc_opaque);
}

using encode_fn = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                               const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_fn get_encode() {
    static encode_fn fn = nullptr;
    if (fn == nullptr) {
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        B200_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres));
        if (sym == nullptr || qres != cudaDriverEntryPointSuccess) { fail(B200SVM_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver"); }
        fn = reinterpret_cast<encode_fn>(sym);
    }
    return fn;
}

// 2-D map over a row-major plane, boxes of 128 bytes x 128 rows (64 fp16 or 128 e4m3 elements), SWIZZLE_128B
void make_map(CUtensorMap *map, void *base, int64_t rows, int64_t ld_elems, bool bytes = false) {
    const cuuint64_t dims[2] = { (cuuint64_t) ld_elems, (cuuint64_t) rows };
    const cuuint64_t strides[1] = { (cuuint64_t) ld_elems * (bytes ? 1 : 2) };
    const cuuint32_t box[2] = { bytes ? 128u : (cuuint32_t) tc::BK, 128u };
    const cuuint32_t estr[2] = { 1u, 1u };
    const CUresult r = get_encode()(map, bytes ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, dims, strides, box, estr,
                                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { fail(B200SVM_ERR_CUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string((int) r)); }
}

// lower block triangle, grouped in super-tiles of GI x GJ block tiles so that one wave of CTAs shares its operands in L2
// grp (optional): local canonical group of every returned tile (canonical_groups(world) equal slices of the full list; a rank owns
// canonical_groups / world consecutive ones — for world = 1, 2, 4, 8 the eight groups are the same sets of tiles)
std::vector<int2> build_tiles(int T_r, int T_c, int rank, int world, std::vector<int> *grp = nullptr) {
    constexpr int GI = 32, GJ = 16;
    std::vector<int2> all;
    for (int sj = 0; sj < T_c; sj += GJ) {
        for (int si = 2 * sj; si < T_r; si += GI) {
            for (int j = sj; j < std::min(sj + GJ, T_c); ++j) {
                for (int i = std::max(si, 2 * j); i < std::min(si + GI, T_r); ++i) { all.push_back(make_int2(i, j)); }
            }
        }
    }
    // contiguous slice of the (uniform-cost) tile list per rank: balanced to within one tile
    const size_t lo = all.size() * (size_t) rank / (size_t) world, hi = all.size() * (size_t) (rank + 1) / (size_t) world;
    if (grp != nullptr) {
        const int G = canonical_groups(world), per_rank = G / world;
        grp->assign(hi - lo, 0);
        int g = rank * per_rank;
        for (size_t t = lo; t < hi; ++t) {
            while (g + 1 < (rank + 1) * per_rank && t >= all.size() * (size_t) (g + 1) / (size_t) G) { ++g; }
            (*grp)[t - lo] = g - rank * per_rank;
        }
    }
    return std::vector<int2>(all.begin() + (std::ptrdiff_t) lo, all.begin() + (std::ptrdiff_t) hi);
}

}  // namespace

bool tc_available(const ctx &c) {
    return c.has_split && c.tc_opaque != nullptr && static_cast<const tc_state *>(c.tc_opaque)->ready;
}

void tc_prepare_split(ctx &c) {
    if (c.dtype != B200SVM_F32 || c.is_csr || c.kernel == B200SVM_LINEAR) { return; }
    const int64_t m = c.m();
    if (m < 1) { return; }
    tc_state &st = state_of(c);
    st.ready = false;
    c.n_pad = (c.n + tc::BN - 1) / tc::BN * tc::BN;
    c.ldh = (c.ldx + tc::BK - 1) / tc::BK * tc::BK;
    // scale: place max |x| in [2^13, 2^14) so that the low plane keeps 11 bits down to |x| = max / 2^13
    st.amax.reserve(sizeof(unsigned int));
    B200_CUDA_CHECK(cudaMemsetAsync(st.amax.p, 0, sizeof(unsigned int), c.stream));
    tc::k_absmax<<<4 * NUM_SMS, 256, 0, c.stream>>>(c.X.as<float>(), c.n * c.ldx, st.amax.as<unsigned int>());
    count_launch(c);
    unsigned int *mb = static_cast<unsigned int *>(c.pinned);
    B200_CUDA_CHECK(cudaMemcpyAsync(mb, st.amax.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, c.stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));
    float amax;
    std::memcpy(&amax, mb, 4);
    float s = 1.f;
    if (amax > 0.f && std::isfinite(amax)) {
        int e = 0;
        std::frexp(amax, &e);  // amax = f * 2^e, f in [0.5, 1)  ->  amax * 2^(14 - e) in [2^13, 2^14)
        s = std::ldexp(1.f, 14 - e);
    }
    c.split_scale = s;
    const size_t plane = (size_t) c.n_pad * c.ldh * sizeof(__half);
    st.fp8x = c.opt_tc_mode != 1;  // 0 (default) / 2 = fp8 cross terms, 1 = three f16 MMAs
    c.Xh_hi.reserve(plane);
    if (st.fp8x) {
        st.P8a.reserve(plane);  // 2 ldh bytes per row
        st.P8b.reserve(plane);
        tc::k_split_fp16<<<8 * NUM_SMS, 256, 0, c.stream>>>(c.X.as<float>(), c.n, c.ldx, s, c.Xh_hi.as<__half>(), nullptr, st.P8a.as<uint8_t>(),
                                                             st.P8b.as<uint8_t>(), c.n_pad, c.ldh);
    } else {
        c.Xh_lo.reserve(plane);
        tc::k_split_fp16<<<8 * NUM_SMS, 256, 0, c.stream>>>(c.X.as<float>(), c.n, c.ldx, s, c.Xh_hi.as<__half>(), c.Xh_lo.as<__half>(), nullptr, nullptr,
                                                             c.n_pad, c.ldh);
    }
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
    make_map(&st.map_hi, c.Xh_hi.p, c.n_pad, c.ldh);
    if (st.fp8x) {
        make_map(&st.map_a8, st.P8a.p, c.n_pad, 2 * c.ldh, true);
        make_map(&st.map_b8, st.P8b.p, c.n_pad, 2 * c.ldh, true);
    } else {
        make_map(&st.map_lo, c.Xh_lo.p, c.n_pad, c.ldh);
    }

    st.T_r = (int) (c.n_pad / tc::BM);
    st.T_c = (int) (c.n_pad / tc::BN);
    st.ldP = c.n_pad;
    std::vector<int> grp;
    const std::vector<int2> tiles = build_tiles(st.T_r, st.T_c, c.rank, c.world, &grp);
    st.groups = canonical_groups(c.world) / c.world;
    st.num_tiles = (int) tiles.size();
    st.tiles.reserve(std::max<size_t>(tiles.size(), 1) * sizeof(int2));
    B200_CUDA_CHECK(cudaMemcpyAsync(st.tiles.p, tiles.data(), tiles.size() * sizeof(int2), cudaMemcpyHostToDevice, c.stream));
    // tile -> local group of this rank (255: another rank's tile), for the reduction of the Gram kernel's per-tile partial vectors
    std::vector<uint8_t> own_map((size_t) st.T_r * (size_t) st.T_c, (uint8_t) 255);
    for (size_t t = 0; t < tiles.size(); ++t) { own_map[(size_t) tiles[t].x * (size_t) st.T_c + (size_t) tiles[t].y] = (uint8_t) grp[t]; }
    st.own.reserve(own_map.size());
    B200_CUDA_CHECK(cudaMemcpyAsync(st.own.p, own_map.data(), own_map.size(), cudaMemcpyHostToDevice, c.stream));
    build_sweep_plan(c, st.sweep, tiles, grp, st.groups, st.T_r, st.T_c);
    st.pk.reserve((size_t) c.n_pad * sizeof(float4));
    st.Pr.reserve((size_t) st.T_c * st.ldP * sizeof(float));
    const size_t pc_rows = (size_t) std::max(st.T_r, st.sweep.slots);  // Gram kernel: one per tile row; sweep: one per (group, run)
    st.Pc.reserve(pc_rows * st.ldP * sizeof(float));
    B200_CUDA_CHECK(cudaMemsetAsync(st.Pr.p, 0, (size_t) st.T_c * st.ldP * sizeof(float), c.stream));
    B200_CUDA_CHECK(cudaMemsetAsync(st.Pc.p, 0, pc_rows * st.ldP * sizeof(float), c.stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(c.stream));  // `tiles` (host vector) must outlive the copy
    static bool attr_set = false;
    if (!attr_set) {
#define B200_TC_ATTR(KN, F8, ST) \
    B200_CUDA_CHECK(cudaFuncSetAttribute(tc::k_gram_tc<KN, F8, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::SMEM_BYTES))
        B200_TC_ATTR(B200SVM_RBF, false, false);
        B200_TC_ATTR(B200SVM_POLYNOMIAL, false, false);
        B200_TC_ATTR(B200SVM_RBF, true, false);
        B200_TC_ATTR(B200SVM_POLYNOMIAL, true, false);
        B200_TC_ATTR(B200SVM_RBF, false, true);
        B200_TC_ATTR(B200SVM_POLYNOMIAL, false, true);
        B200_TC_ATTR(B200SVM_RBF, true, true);
        B200_TC_ATTR(B200SVM_POLYNOMIAL, true, true);
#undef B200_TC_ATTR
        attr_set = true;
    }
    c.has_split = true;
    st.ready = true;
    st.kc_valid = false;
    st.kc_refused = false;
    c.stats.dense_cache_bytes = 0;
}

void tc_invalidate_cache(ctx &c) {
    if (c.tc_opaque != nullptr) {
        tc_state &st = state_of(c);
        st.kc_valid = false;
        st.kc_refused = false;
        c.stats.dense_cache_bytes = 0;
    }
}

namespace {

template <bool STORE>
void launch_gram(ctx &c, tc_state &st, const tc::params &p) {
    const int grid = std::min(NUM_SMS, std::max(st.num_tiles, 1));
    if (st.fp8x) {
        if (c.kernel == B200SVM_RBF) {
            tc::k_gram_tc<B200SVM_RBF, true, STORE><<<grid, tc::NUM_THREADS, tc::SMEM_BYTES, c.stream>>>(st.map_hi, st.map_a8, st.map_b8, p);
        } else {
            tc::k_gram_tc<B200SVM_POLYNOMIAL, true, STORE><<<grid, tc::NUM_THREADS, tc::SMEM_BYTES, c.stream>>>(st.map_hi, st.map_a8, st.map_b8, p);
        }
    } else if (c.kernel == B200SVM_RBF) {
        tc::k_gram_tc<B200SVM_RBF, false, STORE><<<grid, tc::NUM_THREADS, tc::SMEM_BYTES, c.stream>>>(st.map_hi, st.map_lo, st.map_lo, p);
    } else {
        tc::k_gram_tc<B200SVM_POLYNOMIAL, false, STORE><<<grid, tc::NUM_THREADS, tc::SMEM_BYTES, c.stream>>>(st.map_hi, st.map_lo, st.map_lo, p);
    }
    count_launch(c);
    B200_CUDA_CHECK(cudaGetLastError());
}

}  // namespace

void *tc_pack_buffer(ctx &c) { return tc_available(c) ? state_of(c).pk.p : nullptr; }

// out[i] = sum_j (k(x_i, x_j) + QA_cost - q_i - q_j) d_j  over the tiles owned by this rank (all tiles when world == 1);
// push != nullptr: the sums go into slot `rank` of every peer's exchange buffer instead (the reduction kernel is the producer)
void tc_kernel_matvec(ctx &c, const float *d, float *out, bool packed, const xchg_push *push, const float *done) {
    if (!tc_available(c)) { fail(B200SVM_ERR_INVALID, "tcgen05 matvec path requested but the fp16 operand planes are not prepared"); }
    tc_state &st = state_of(c);
    const int64_t m = c.m();
    constexpr float LOG2E = 1.4426950408889634f;
    const float s2 = c.split_scale * c.split_scale;
    tc::params p{};
    p.tiles = st.tiles.as<int2>();
    p.num_tiles = st.num_tiles;
    p.num_k_blocks = (int) (c.ldh / tc::BK);
    p.pk = st.pk.as<float4>();
    p.Pr = st.Pr.as<float>();
    p.Pc = st.Pc.as<float>();
    p.ldP = st.ldP;
    p.coef0 = (float) c.coef0;
    p.qa = (float) c.QA_cost;
    p.degree = c.degree;
    p.kernel = c.kernel;
    p.Kc = nullptr;
    p.done = done;
    float cscale = 0.f;
    if (c.kernel == B200SVM_RBF) {
        p.w = (float) (2.0 * c.gamma * (double) LOG2E / (double) s2);
        cscale = (float) (-c.gamma * (double) LOG2E);
    } else {
        p.w = (float) (c.gamma / (double) s2);
    }
    if (!packed) {  // (inside the CG loop the direction update keeps the packed operand current: cg.cu)
        tc::k_pack<<<(unsigned) ((c.n_pad + 255) / 256), 256, 0, c.stream>>>(c.norms.as<float>(), c.q.as<float>(), d, m, c.n_pad, cscale, st.pk.as<float4>());
        count_launch(c);
    }
    // The kernel values depend only on the data: when the tiles of this rank fit in HBM they are stored once (one pass of the
    // tcgen05 kernel in STORE mode, which also reduces against d: the assembling matvec needs no sweep) and every later matvec is
    // an HBM-bound sweep over them.
    bool assembled_now = false;
    if (c.opt_dense_cache != 1 && !st.kc_valid && !st.kc_refused && st.num_tiles > 0) {
        const size_t need = (size_t) st.num_tiles * tc::BM * tc::BN * sizeof(float);
        size_t free_b = 0, total_b = 0;
        B200_CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
        free_b += st.Kc.bytes;
        const double budget = c.opt_sparse_cache_mb > 0 ? (double) c.opt_sparse_cache_mb * 1048576.0 : 0.85 * (double) free_b;
        if ((double) need > budget) {
            B200_REQUIRE(c.opt_dense_cache != 2, "the cached kernel-matrix tiles do not fit the device memory budget");
            st.kc_refused = true;
        } else {
            st.Kc.reserve(need);
            tc::params ps = p;
            ps.Kc = st.Kc.as<float>();
            ps.done = nullptr;  // never skipped: the stored tiles outlive the solve
            if (c.cb_pending) { sync_stream(c); }
            B200_CUDA_CHECK(cudaEventRecord(c.cb_ev[0], c.stream));
            launch_gram<true>(c, st, ps);
            B200_CUDA_CHECK(cudaEventRecord(c.cb_ev[1], c.stream));
            c.cb_pending = true;
            st.kc_valid = true;
            assembled_now = true;
            c.stats.dense_cache_bytes = (int64_t) need;
        }
    }
    if (c.opt_dense_cache == 1) { st.kc_valid = false; }
    tc::reduce_dst dst{};
    dst.out = out;
    dst.remote = push != nullptr ? 1 : 0;
    if (push != nullptr) { dst.push = *push; }
    const bool swept = st.kc_valid && !assembled_now;  // (the STORE pass has already filled Pr / Pc for this d)
    if (swept) {
        tc::sweep_params sp{};
        sp.Kc = st.Kc.as<float>();
        sp.list = st.sweep.list.as<int4>();
        sp.num_tiles = st.num_tiles;
        sp.nchunks = st.sweep.nchunks;
        sp.chunks = st.sweep.chunks.as<int4>();
        sp.pk = p.pk;
        sp.Pr = p.Pr;
        sp.Pc = p.Pc;
        sp.ldP = p.ldP;
        sp.qa = p.qa;
        sp.done = done;
        tc::k_tile_sweep<true><<<st.sweep.nchunks, 256, 0, c.stream>>>(sp);
        tc::k_sweep_reduce<float><<<tc::reduce_grid(m, st.sweep.groups), 256, 0, c.stream>>>(p.Pr, p.Pc, p.ldP, m, st.T_c, st.sweep.jptr.as<int>(), st.sweep.jlist.as<int>(),
                                                                              st.sweep.nruns.as<int>(), st.sweep.slot_base.as<int>(), st.sweep.groups, dst, done);
    } else {
        if (!st.kc_valid) { launch_gram<false>(c, st, p); }
        tc::k_tc_reduce<float><<<tc::reduce_grid(m, st.groups), 256, 0, c.stream>>>(p.Pr, p.Pc, p.ldP, m, st.T_r, st.T_c, st.own.as<uint8_t>(), st.groups, dst, done);
    }
    count_launch(c, swept ? 2 : 1);
    B200_CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------------------------------
// Symmetric tile store for the sparse path: when S = f(X X^T) of a CSR data set is densely filled, it is kept as
// dense 128 x 256 tiles of the lower block triangle (4 bytes per stored entry, no indices, half the matrix) and the CG matvec is
// the same HBM-bound sweep as for dense data (k_tile_sweep<false>: un-centred, the rank-1 part is added by the O(m) epilogue).
// Row blocks are dealt to the ranks in contiguous ranges of equal tile count; storage index = position in the row-block-major
// list, so sparse.cu can address tile (I, J) of a row without a table: tile_store_cum(I) - tile_store_cum(I_lo) + J.
// ---------------------------------------------------------------------------------------------------------------------
struct tile_store {
    sweep_plan sweep;
    dbuf pk, Pr, Pc, Kc;
    size_t esize{ 4 };   // 4: float tiles (sparse S), 8: double tiles (dense double kernel matrix)
    int T_r{ 0 }, T_c{ 0 };
    int I_lo{ 0 }, I_hi{ 0 };
    int64_t n_pad{ 0 }, ldP{ 0 };
    int num_tiles{ 0 };
};

namespace tc {
// pk[i] = {0, 0, w_i, 0} for i < m, zeros above
__global__ void __launch_bounds__(256) k_pack_plain(const float *__restrict__ w, int64_t m, int64_t n_pad, float4 *__restrict__ pk) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pad) { return; }
    pk[i] = make_float4(0.f, 0.f, i < m ? w[i] : 0.f, 0.f);
}
}  // namespace tc

int64_t tile_store_cum(int64_t I) {  // tiles (I', J), 2 J <= I', of the row blocks I' < I
    const int64_t k = I >> 1;
    return (I & 1) ? (k + 1) * (k + 1) : k * (k + 1);
}

// row-block boundary of the `part`-th of `parts` equal shares of the tiles (by cumulative tile count)
static int tile_store_bound(int T_r, int part, int parts) {
    if (part <= 0) { return 0; }
    if (part >= parts) { return T_r; }
    const int64_t want = tile_store_cum(T_r) * part / parts;
    int I = 0;
    while (I < T_r && tile_store_cum(I) < want) { ++I; }
    return I;
}

// rows of a rank = its canonical_groups / world consecutive groups of row blocks (the same eight groups for world = 1, 2, 4, 8)
void tile_store_range(int64_t m, int rank, int world, int *I_lo, int *I_hi) {
    const int64_t n_pad = (m + tc::BN - 1) / tc::BN * tc::BN;
    const int T_r = (int) (n_pad / tc::BM);
    const int G = canonical_groups(world), per_rank = G / world;
    *I_lo = tile_store_bound(T_r, rank * per_rank, G);
    *I_hi = tile_store_bound(T_r, (rank + 1) * per_rank, G);
}

tile_store *tile_store_create(ctx &c, int64_t m, size_t esize) {
    tile_store *ts = new tile_store();
    ts->esize = esize;
    ts->n_pad = (m + tc::BN - 1) / tc::BN * tc::BN;
    ts->T_r = (int) (ts->n_pad / tc::BM);
    ts->T_c = (int) (ts->n_pad / tc::BN);
    ts->ldP = ts->n_pad;
    tile_store_range(m, c.rank, c.world, &ts->I_lo, &ts->I_hi);
    std::vector<int2> tiles;
    std::vector<int> grp;
    const int G = canonical_groups(c.world), per_rank = G / c.world;
    for (int g = 0; g < per_rank; ++g) {
        const int Ia = tile_store_bound(ts->T_r, c.rank * per_rank + g, G), Ib = tile_store_bound(ts->T_r, c.rank * per_rank + g + 1, G);
        for (int I = Ia; I < Ib; ++I) {
            for (int J = 0; 2 * J <= I; ++J) {
                tiles.push_back(make_int2(I, J));
                grp.push_back(g);
            }
        }
    }
    ts->num_tiles = (int) tiles.size();
    try {
        build_sweep_plan(c, ts->sweep, tiles, grp, per_rank, ts->T_r, ts->T_c);
        ts->pk.reserve((size_t) ts->n_pad * 16);  // float4 {-, -, w, -} or double2 {q, w}
        ts->Pr.reserve((size_t) ts->T_c * ts->ldP * esize);
        ts->Pc.reserve((size_t) ts->sweep.slots * ts->ldP * esize);
        B200_CUDA_CHECK(cudaMemsetAsync(ts->Pr.p, 0, ts->Pr.bytes, c.stream));
        B200_CUDA_CHECK(cudaMemsetAsync(ts->Pc.p, 0, ts->Pc.bytes, c.stream));
    } catch (...) {
        delete ts;
        throw;
    }
    return ts;
}

void tile_store_destroy(tile_store *ts) { delete ts; }
int64_t tile_store_tiles(const tile_store *ts) { return ts->num_tiles; }
size_t tile_store_bytes(const tile_store *ts) { return (size_t) ts->num_tiles * tc::BM * tc::BN * ts->esize; }
void tile_store_rows(const tile_store *ts, int64_t *row_lo, int64_t *row_hi, int64_t *first_tile) {
    *row_lo = (int64_t) ts->I_lo * tc::BM;
    *row_hi = (int64_t) ts->I_hi * tc::BM;
    *first_tile = tile_store_cum(ts->I_lo);
}
size_t tile_store_allocated(const tile_store *ts) { return ts->Kc.bytes; }  // bytes already held for the tiles (reused by the next assembly)
void *tile_store_alloc(tile_store *ts) {
    ts->Kc.reserve(std::max<size_t>(tile_store_bytes(ts), 16));
    return ts->Kc.p;
}

// out[i] = sum_j S_ij w_j over the tiles of this rank (a full-length partial vector when world > 1)
void tile_store_matvec(ctx &c, tile_store *ts, const float *w, int64_t m, float *out, const float *done) {
    tc::k_pack_plain<<<(unsigned) ((ts->n_pad + 255) / 256), 256, 0, c.stream>>>(w, m, ts->n_pad, ts->pk.as<float4>());
    tc::sweep_params sp{};
    sp.Kc = ts->Kc.as<float>();
    sp.list = ts->sweep.list.as<int4>();
    sp.num_tiles = ts->num_tiles;
    sp.nchunks = ts->sweep.nchunks;
    sp.chunks = ts->sweep.chunks.as<int4>();
    sp.pk = ts->pk.as<float4>();
    sp.Pr = ts->Pr.as<float>();
    sp.Pc = ts->Pc.as<float>();
    sp.ldP = ts->ldP;
    sp.qa = 0.f;
    sp.done = done;
    if (ts->num_tiles > 0) { tc::k_tile_sweep<false><<<ts->sweep.nchunks, 256, 0, c.stream>>>(sp); }
    tc::reduce_dst dst{};
    dst.out = out;
    dst.remote = 0;
    tc::k_sweep_reduce<float><<<tc::reduce_grid(m, ts->sweep.groups), 256, 0, c.stream>>>(sp.Pr, sp.Pc, sp.ldP, m, ts->T_c, ts->sweep.jptr.as<int>(), ts->sweep.jlist.as<int>(),
                                                                          ts->sweep.nruns.as<int>(), ts->sweep.slot_base.as<int>(), ts->sweep.groups, dst, done);
    count_launch(c, 3);
    B200_CUDA_CHECK(cudaGetLastError());
}

// out[i] = sum_j (K_ij + QA_cost - q_i - q_j) w_j over the double tiles of this rank; push != nullptr: straight into the peers
void tile_store_matvec_f64(ctx &c, tile_store *ts, const double *q, const double *w, double qa, int64_t m, double *out, const xchg_push *push,
                           const double *done) {
    tc::k_pack_f64<<<(unsigned) ((ts->n_pad + 255) / 256), 256, 0, c.stream>>>(q, w, m, ts->n_pad, ts->pk.as<double2>());
    tc::sweep_params_f64 sp{};
    sp.Kc = ts->Kc.as<double>();
    sp.list = ts->sweep.list.as<int4>();
    sp.num_tiles = ts->num_tiles;
    sp.nchunks = ts->sweep.nchunks;
    sp.chunks = ts->sweep.chunks.as<int4>();
    sp.pk = ts->pk.as<double2>();
    sp.Pr = ts->Pr.as<double>();
    sp.Pc = ts->Pc.as<double>();
    sp.ldP = ts->ldP;
    sp.qa = qa;
    sp.done = done;
    if (ts->num_tiles > 0) { tc::k_tile_sweep_f64<<<std::min(ts->sweep.nchunks, 2 * NUM_SMS), 512, 0, c.stream>>>(sp); }
    tc::reduce_dst_t<double> dst{};
    dst.out = out;
    dst.remote = push != nullptr ? 1 : 0;
    if (push != nullptr) { dst.push = *push; }
    tc::k_sweep_reduce<double><<<tc::reduce_grid(m, ts->sweep.groups), 256, 0, c.stream>>>(sp.Pr, sp.Pc, sp.ldP, m, ts->T_c, ts->sweep.jptr.as<int>(), ts->sweep.jlist.as<int>(),
                                                                                ts->sweep.nruns.as<int>(), ts->sweep.slot_base.as<int>(), ts->sweep.groups, dst, done);
    count_launch(c, 3);
    B200_CUDA_CHECK(cudaGetLastError());
}

}  // namespace b200

extern "C" int b200svm_partition_tiles(int64_t n, int rank, int world, int64_t *first, int64_t *count, int64_t *total) {
    if (first == nullptr || count == nullptr || total == nullptr || world < 1 || rank < 0 || rank >= world || n < 1) { return B200SVM_ERR_INVALID; }
    const int64_t n_pad = (n + b200::tc::BN - 1) / b200::tc::BN * b200::tc::BN;
    const int T_r = (int) (n_pad / b200::tc::BM), T_c = (int) (n_pad / b200::tc::BN);
    const size_t all = b200::build_tiles(T_r, T_c, 0, 1).size();
    const size_t lo = all * (size_t) rank / (size_t) world, hi = all * (size_t) (rank + 1) / (size_t) world;
    *first = (int64_t) lo;
    *count = (int64_t) (hi - lo);
    *total = (int64_t) all;
    return B200SVM_OK;
}

extern "C" int b200svm_canonical_groups(int world) { return world >= 1 ? b200::canonical_groups(world) : 0; }

extern "C" int b200svm_partition_tile_rows(int64_t m, int rank, int world, int64_t *row_lo, int64_t *row_hi) {
    if (row_lo == nullptr || row_hi == nullptr || world < 1 || rank < 0 || rank >= world || m < 1) { return B200SVM_ERR_INVALID; }
    int I_lo = 0, I_hi = 0;
    b200::tile_store_range(m, rank, world, &I_lo, &I_hi);
    *row_lo = (int64_t) I_lo * b200::tc::BM;
    *row_hi = (int64_t) I_hi * b200::tc::BM;
    return B200SVM_OK;
}
