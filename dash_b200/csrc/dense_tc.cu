/*
 * dense_tc.cu -- garbled dense layer on the 5th-generation tensor cores.
 *
 *   out_j[m][n] = ( sum_k Wq_j[m][k] * in_j[k][n] + cst_j[m][d(n)] ) mod p_j
 *
 * for every CRT modulus j in one launch.  The digit planes are already the
 * row-major K x N operand (N = 4*nc4_j*Bp bytes per element row), so they are
 * consumed as they are: TMA (cp.async.bulk.tensor, SWIZZLE_128B) stages
 * 128 x 128-byte tiles of the weights (K-major) and of the plane (N-major =
 * "MN-major" B operand) in shared memory, one elected thread issues
 * tcgen05.mma.kind::i8 (u8 x u8 -> s32, accumulator in TMEM), four epilogue
 * warps read the accumulator back with tcgen05.ld, add the constant label,
 * reduce mod p_j and store bytes.  Exact while K * 255^2 < 2^31 (checked by
 * the caller).  Replaces MatMulModAddZero (misc/cuda_util.h:95-114) /
 * LabelTensor::matvecmul (garbling/label_tensor.h:783-818); same
 * reformulation of the zero-weight rule as k_linear (kernels.cu).
 *
 * Warp roles (192 threads): warp 0 = TMA producer, warp 1 = MMA issuer +
 * TMEM allocation, warps 2..5 = epilogue (TMEM sub-partition = warp % 4).
 * A shared-memory ring (full/empty mbarriers) decouples them.
 *
 * Convolution (CONV = true) is the same GEMM per output pixel: M = filters,
 * K = taps (channel, filter row, filter column), N = the pixel's plane row.
 * The im2col gather is done by TMA itself: the input plane is described as a
 * 4-D tensor {plane bytes, column, row, channel} and one box {128, f_w, f_h,
 * Cb} lands in shared memory as Cb*f_h*f_w consecutive 128-byte K rows --
 * exactly the N-major B tile of the dense case.  The weights are laid out to
 * match (Cb channels per 128-column K block, zero padded), so the rows of a
 * K block that the box does not fill meet zero weights.  Replaces
 * Conv2dModZero (misc/cuda_util.h:209-268) / LabelTensor::conv2d
 * (garbling/label_tensor.h:920-1002) for layers wide enough to fill a tile.
 */
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

#include "kernels.h"

namespace dashb {

using gc::ModInfo;

constexpr int TC_BM = 128, TC_BN = 128, TC_BK = 128; /* u8: bytes == elements */
constexpr int TC_STAGES_DENSE = 4; /* long K loops: deep ring, one CTA per SM */
constexpr int TC_STAGES_CONV = 2;  /* few K blocks per tile: several CTAs per SM overlap MMA and epilogue */
/* column sub-tiles of 128 per CTA: the weight tile staged once feeds NB MMAs.
 * The kernel is bound by L2 -> shared-memory operand traffic (profiles/
 * r1_dense_tc_vs_dp4a.md), NB = 2 moves 48 KB instead of 64 KB per two tiles. */
#ifndef DASHB_TC_NB
#define DASHB_TC_NB 2
#endif
constexpr int TC_NB = DASHB_TC_NB;
constexpr int TC_THREADS = 192;
__host__ __device__ constexpr uint32_t tc_stage_bytes(int nb) { return TC_BM * TC_BK + (uint32_t)nb * TC_BK * TC_BN; }
__host__ __device__ constexpr uint32_t tc_smem_bytes(int stages, int nb) { return (uint32_t)stages * tc_stage_bytes(nb) + 1024; }

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
/* bounded wait: a protocol error traps instead of hanging the GPU */
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t it = 0; it < (1u << 26); ++it) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) return;
    }
    __trap();
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
/* shared-memory matrix descriptor: SWIZZLE_128B, version 1 (sm_100) */
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

template <bool CONV, int TC_STAGES, int NB>
__global__ void __launch_bounds__(TC_THREADS, 1)
k_dense_tc(const __grid_constant__ DenseTcMaps maps, const __grid_constant__ DenseTcParams P) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar_full[TC_STAGES], bar_empty[TC_STAGES], bar_acc;
    __shared__ uint32_t tmem_base_s;
    const int j = blockIdx.z;
    const ModInfo m = P.mod[j];
    const uint32_t nc4 = (m.nc + 3) >> 2;
    const uint32_t Nj = nc4 * 4 * P.Bp;
    constexpr uint32_t TC_STAGE_BYTES = tc_stage_bytes(NB);
    const uint32_t n0 = blockIdx.x * (TC_BN * NB);
    if (n0 >= Nj) return; /* uniform per CTA: planes of small moduli have fewer columns */
    /* column sub-tiles that start inside the plane */
    const uint32_t nb_act = min((uint32_t)NB, (Nj - n0 + TC_BN - 1) / TC_BN);
    /* conv: blockIdx.y = filter block * npix + output pixel */
    /* (a launch covers the pixels [p0, p1): all of them, or this rank's share) */
    const uint32_t pix = CONV ? P.p0 + blockIdx.y % (P.p1 - P.p0) : 0u;
    const uint32_t m0 = P.v0 + (CONV ? blockIdx.y / (P.p1 - P.p0) : blockIdx.y) * TC_BM;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t nkb = P.nkb;

    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(TC_BN * NB));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(&bar_full[s], 1);
            mbar_init(&bar_empty[s], 1);
        }
        mbar_init(&bar_acc, 1);
        asm volatile("fence.mbarrier_init.release.cluster;");
        asm volatile("fence.proxy.async.shared::cta;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        /* ===== TMA producer ===== */
        if (lane == 0) {
            for (uint32_t kb = 0; kb < nkb; ++kb) {
                const uint32_t s = kb % TC_STAGES, round = kb / TC_STAGES;
                if (round > 0) mbar_wait(&bar_empty[s], (round - 1) & 1);
                uint8_t* sA = smem + s * TC_STAGE_BYTES;
                uint8_t* sB = sA + TC_BM * TC_BK;
                mbar_expect_tx(&bar_full[s], TC_BM * TC_BK + nb_act * P.b_box_bytes);
                tma_load_2d(sA, reinterpret_cast<const CUtensorMap*>(&maps.a[j]), &bar_full[s], (int)(kb * TC_BK), (int)m0);  /* {k, m} */
                for (uint32_t h = 0; h < nb_act; ++h) {
                    uint8_t* sBh = sB + h * (TC_BK * TC_BN);
                    const int nh = (int)(n0 + h * TC_BN);
                    if (CONV)
                        tma_load_4d(sBh, reinterpret_cast<const CUtensorMap*>(&maps.b[j]), &bar_full[s], nh,
                                    (int)((pix % P.ow) * P.sx), (int)((pix / P.ow) * P.sy), (int)(kb * P.cb)); /* {n, x, y, c} */
                    else
                        tma_load_2d(sBh, reinterpret_cast<const CUtensorMap*>(&maps.b[j]), &bar_full[s], nh, (int)(kb * TC_BK));  /* {n, k} */
                }
            }
        }
    } else if (warp == 1) {
        /* ===== MMA issuer ===== */
        if (lane == 0) {
            /* S32 accumulator, U8 x U8, A K-major, B MN-major, M = 128, N = 128 */
            const uint32_t idesc = (2u << 4) | (1u << 16) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
            for (uint32_t kb = 0; kb < nkb; ++kb) {
                const uint32_t s = kb % TC_STAGES, round = kb / TC_STAGES;
                mbar_wait(&bar_full[s], round & 1);
                asm volatile("tcgen05.fence::after_thread_sync;");
                const uint32_t aA = smem_u32(smem + s * TC_STAGE_BYTES);
                const uint32_t aB = aA + TC_BM * TC_BK;
                for (uint32_t h = 0; h < nb_act; ++h) {
#pragma unroll
                    for (int k4 = 0; k4 < TC_BK / 32; ++k4) {
                        const uint64_t da = umma_desc(aA + k4 * 32, 16, 1024);                                  /* +32 B inside the swizzled row */
                        const uint64_t db = umma_desc(aB + h * (TC_BK * TC_BN) + k4 * 32 * TC_BN, 16, 1024);    /* +32 k-rows of 128 B */
                        const uint32_t acc = (kb | (uint32_t)k4) ? 1u : 0u;
                        asm volatile(
                            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}" ::"r"(tmem_base + h * TC_BN),
                            "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(0), "r"(0), "r"(0), "r"(0)
                            : "memory");
                    }
                }
                /* frees the stage when the MMAs that read it are done */
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                 smem_u32(&bar_empty[s]))
                             : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar_acc))
                         : "memory");
        }
    } else {
        /* ===== epilogue: TMEM -> registers -> (+cst) mod q -> bytes ===== */
        mbar_wait(&bar_acc, 0);
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t sub = (uint32_t)warp & 3u; /* TMEM lanes 32*sub .. 32*sub+31 */
        const uint32_t row = m0 + sub * 32 + lane;
        const bool row_ok = row < P.v1;
        const uint32_t* cst = reinterpret_cast<const uint32_t*>(P.cst[j]) + (uint64_t)(row_ok ? row : 0) * nc4;
        uint8_t* orow = P.out[j] + ((uint64_t)(row_ok ? row : 0) * (CONV ? P.npix : 1u) + pix) * Nj + n0;
        for (int c0 = 0; c0 < TC_BN * (int)nb_act; c0 += 32) {
            uint32_t v[32];
            const uint32_t taddr = tmem_base + ((sub * 32u) << 16) + (uint32_t)c0;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                  "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                  "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                  "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (row_ok && n0 + c0 < Nj) {
                uint32_t w[8];
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    /* 4 columns = digits 4*d4..4*d4+3 of one inference */
                    const uint32_t d4 = (n0 + c0 + 4 * g) / (4 * P.Bp);
                    const uint32_t cw = __ldg(cst + d4);
                    uint32_t o = 0;
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        uint32_t x = v[4 * g + t] + ((cw >> (8 * t)) & 0xff);
                        x -= gc::mulhi32(x, m.mq) * m.q; /* reciprocal over-estimates by at most one */
                        if ((int32_t)x < 0) x += m.q;
                        o |= x << (8 * t);
                    }
                    w[g] = o;
                }
                /* Nj is a multiple of 64, tiles of 32 columns never straddle it */
                *reinterpret_cast<uint4*>(orow + c0) = make_uint4(w[0], w[1], w[2], w[3]);
                *reinterpret_cast<uint4*>(orow + c0 + 16) = make_uint4(w[4], w[5], w[6], w[7]);
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TC_BN * NB));
}

/*
 * Convolution with the FILTERS on the N dimension of the MMA: D[plane byte]
 * [filter] = sum_k plane[k][byte] * W[filter][k].  The A operand (M = 128) is
 * the im2col'd plane tile exactly as the kernel above stages it (now an
 * MN-major A), the B operand the weight rows (K-major).  An MMA then costs
 * N/2 = filters/2 cycles instead of 64, every TMEM lane holds a result (a
 * lane is a plane byte), so all four epilogue warps work whatever the filter
 * count -- with the filters on M, 32 filters fill a quarter of the tile and
 * keep one epilogue warp busy.  For 32 | filters <= 128.
 */
template <int TC_STAGES>
__global__ void __launch_bounds__(TC_THREADS, 1)
k_conv_tc_n(const __grid_constant__ DenseTcMaps maps, const __grid_constant__ DenseTcParams P) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar_full[TC_STAGES], bar_empty[TC_STAGES], bar_acc;
    __shared__ uint32_t tmem_base_s;
    constexpr uint32_t STAGE = TC_BK * TC_BN + TC_BM * TC_BK; /* plane tile, then weight tile */
    const int j = blockIdx.z;
    const ModInfo m = P.mod[j];
    const uint32_t nc4 = (m.nc + 3) >> 2;
    const uint32_t Nj = nc4 * 4 * P.Bp;
    const uint32_t n0 = blockIdx.x * TC_BN;
    if (n0 >= Nj) return;
    const uint32_t pix = P.p0 + blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t nkb = P.nkb, npad = P.npad;
    const uint32_t tmem_cols = npad <= 32 ? 32u : npad <= 64 ? 64u : 128u; /* a power of two >= 32 */

    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(&bar_full[s], 1);
            mbar_init(&bar_empty[s], 1);
        }
        mbar_init(&bar_acc, 1);
        asm volatile("fence.mbarrier_init.release.cluster;");
        asm volatile("fence.proxy.async.shared::cta;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        if (lane == 0) {
            for (uint32_t kb = 0; kb < nkb; ++kb) {
                const uint32_t s = kb % TC_STAGES, round = kb / TC_STAGES;
                if (round > 0) mbar_wait(&bar_empty[s], (round - 1) & 1);
                uint8_t* sP = smem + s * STAGE;
                uint8_t* sW = sP + TC_BK * TC_BN;
                mbar_expect_tx(&bar_full[s], TC_BM * TC_BK + P.b_box_bytes);
                tma_load_4d(sP, reinterpret_cast<const CUtensorMap*>(&maps.b[j]), &bar_full[s], (int)n0, (int)((pix % P.ow) * P.sx),
                            (int)((pix / P.ow) * P.sy), (int)(kb * P.cb));
                tma_load_2d(sW, reinterpret_cast<const CUtensorMap*>(&maps.a[j]), &bar_full[s], (int)(kb * TC_BK), 0); /* all filter rows */
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            /* S32 accumulator, U8 x U8, A MN-major (bit 15), B K-major, M = 128, N = npad */
            const uint32_t idesc = (2u << 4) | (1u << 15) | ((npad >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
            for (uint32_t kb = 0; kb < nkb; ++kb) {
                const uint32_t s = kb % TC_STAGES, round = kb / TC_STAGES;
                mbar_wait(&bar_full[s], round & 1);
                asm volatile("tcgen05.fence::after_thread_sync;");
                const uint32_t aP = smem_u32(smem + s * STAGE);
                const uint32_t aW = aP + TC_BK * TC_BN;
#pragma unroll
                for (int k4 = 0; k4 < TC_BK / 32; ++k4) {
                    const uint64_t da = umma_desc(aP + k4 * 32 * TC_BN, 16, 1024); /* +32 k-rows of 128 plane bytes */
                    const uint64_t db = umma_desc(aW + k4 * 32, 16, 1024);         /* +32 B inside the swizzled weight row */
                    const uint32_t acc = (kb | (uint32_t)k4) ? 1u : 0u;
                    asm volatile(
                        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}" ::"r"(tmem_base),
                        "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(0), "r"(0), "r"(0), "r"(0)
                        : "memory");
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                 smem_u32(&bar_empty[s]))
                             : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar_acc))
                         : "memory");
        }
    } else {
        /* epilogue: TMEM lane = plane byte, column = filter.  The reduced
         * bytes are transposed through shared memory (the operand stages are
         * free once the accumulator is complete) so that the global stores
         * are 16 bytes per thread along a filter's row, and the constant
         * label bytes come from a small shared copy instead of one global
         * load per output. */
        mbar_wait(&bar_acc, 0);
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t sub = (uint32_t)warp & 3u;
        const uint32_t te = threadIdx.x - 64; /* 0..127 among the epilogue threads */
        const uint32_t l128 = sub * 32 + lane;
        const uint32_t n = n0 + l128;
        const uint32_t ncp = nc4 * 4;
        const uint32_t d4_first = n0 / (4 * P.Bp);
        uint8_t* s_out = smem;                                        /* [npad][128] */
        uint32_t* s_cst = reinterpret_cast<uint32_t*>(smem + 128 * 128); /* [npad][4] words of 4 digits */
        for (uint32_t i = te; i < npad * 4; i += 128) {
            const uint32_t f = i >> 2, d4 = d4_first + (i & 3u);
            s_cst[i] = d4 < nc4 ? __ldg(reinterpret_cast<const uint32_t*>(P.cst[j] + (uint64_t)f * ncp) + d4) : 0u;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const uint32_t q = (n < Nj ? n / (4 * P.Bp) : d4_first) - d4_first; /* < 4: a 128-byte chunk spans at most 3 digit words */
        const uint32_t sh = 8 * (n & 3u);
        for (uint32_t c0 = 0; c0 < npad; c0 += 32) {
            uint32_t v[32];
            const uint32_t taddr = tmem_base + ((sub * 32u) << 16) + c0;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                  "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                  "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                  "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const uint32_t f = c0 + i; /* filter */
                uint32_t x = v[i] + ((s_cst[f * 4 + q] >> sh) & 0xffu);
                x -= gc::mulhi32(x, m.mq) * m.q;
                if ((int32_t)x < 0) x += m.q;
                s_out[f * 128 + l128] = (uint8_t)x;
            }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        /* 8 threads x 16 bytes per filter row, 16 rows per pass */
        const uint32_t col16 = (te & 7u) * 16;
        if (n0 + col16 < Nj) { /* Nj is a multiple of 64 */
            for (uint32_t f = te >> 3; f < npad; f += 16) {
                if (f < P.v0 || f >= P.v1) continue;
                const uint4 w = *reinterpret_cast<const uint4*>(s_out + f * 128 + col16);
                *reinterpret_cast<uint4*>(P.out[j] + ((uint64_t)f * P.npix + pix) * Nj + n0 + col16) = w;
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols));
}

/*
 * Convolution, stride 1 in x, a STRIP of horizontally adjacent output pixels
 * per CTA (filters on N as in k_conv_tc_n).  The kernels above give every
 * output pixel a CTA of its own, so an input byte is fetched f_w * f_h times
 * and the weight tile once per pixel: they are bound by L2 -> shared-memory
 * traffic (GNNP 32->32 3x3: 5.5 TB/s for 3 % of the int8 tensor peak).  Here
 * the K order is (filter column j; channel c, filter row i): ONE input column
 * x contributes the rows (c, i) -- a TMA box {128 plane bytes, 1 column, f_h
 * rows, cblk channels}, cblk * f_h a multiple of the MMA's K = 32 -- and
 * output pixel p uses the columns p .. p + f_w - 1, each of them a contiguous,
 * uniformly strided K block of the MN-major A operand.  A strip of P pixels
 * loads P + f_w - 1 columns instead of P * f_w and the weights once: 3x less
 * operand traffic for 3x3 filters at P = 4.  Accumulators: P * npad TMEM
 * columns.  MMAs of pixel p are issued as soon as its last column has landed
 * (one mbarrier per column slot), so the loads of the later columns overlap
 * the MMAs of the earlier pixels; two CTAs per SM overlap the rest.
 * Same results as Conv2dModZero / LabelTensor::conv2d (see k_dense_tc).
 */
constexpr int ST_MAXCOL = 8; /* column slots: P + f_w - 1 */
__global__ void __launch_bounds__(TC_THREADS, 1)
k_conv_tc_strip(const __grid_constant__ DenseTcMaps maps, const __grid_constant__ DenseTcParams P) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar_col[ST_MAXCOL], bar_w, bar_free, bar_acc;
    __shared__ uint32_t tmem_base_s;
    const int j = blockIdx.z;
    const ModInfo m = P.mod[j];
    const uint32_t nc4 = (m.nc + 3) >> 2;
    const uint32_t Nj = nc4 * 4 * P.Bp;
    const uint32_t n0 = blockIdx.x * TC_BN;
    if (n0 >= Nj) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t npad = P.npad, S = P.strip, R = P.st_rows, fw = P.st_fw, ncb = P.st_ncb;
    const uint32_t ncol = S + fw - 1;
    /* blockIdx.y = (output row - first row) * strips per row + strip */
    const uint32_t spr = (P.ow + S - 1) / S;
    const uint32_t oy = P.p0 / P.ow + blockIdx.y / spr;
    const uint32_t x0 = (blockIdx.y % spr) * S;
    const uint32_t col_bytes = R * 128u;
    uint8_t* sCol = smem;                      /* [ncol][R][128] */
    uint8_t* sW = smem + ncol * col_bytes;     /* [wboxes][128][128] */
    const uint32_t tmem_cols = S * npad <= 32 ? 32u : S * npad <= 64 ? 64u : S * npad <= 128 ? 128u : 256u;

    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (threadIdx.x == 0) {
        for (int c = 0; c < ST_MAXCOL; ++c) mbar_init(&bar_col[c], 1);
        mbar_init(&bar_w, 1);
        mbar_init(&bar_free, 1);
        mbar_init(&bar_acc, 1);
        asm volatile("fence.mbarrier_init.release.cluster;");
        asm volatile("fence.proxy.async.shared::cta;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        if (lane == 0) {
            for (uint32_t cb = 0; cb < ncb; ++cb) {
                /* one buffer: the next channel block's operands replace this one's once its MMAs have read them */
                if (cb > 0) mbar_wait(&bar_free, (cb - 1) & 1);
                mbar_expect_tx(&bar_w, P.st_wboxes * 128u * 128u);
                for (uint32_t b = 0; b < P.st_wboxes; ++b)
                    tma_load_2d(sW + b * 128u * 128u, reinterpret_cast<const CUtensorMap*>(&maps.a[j]), &bar_w, 0,
                                (int)(cb * fw * npad + b * 128u));
                for (uint32_t c = 0; c < ncol; ++c) {
                    mbar_expect_tx(&bar_col[c], col_bytes);
                    tma_load_4d(sCol + c * col_bytes, reinterpret_cast<const CUtensorMap*>(&maps.b[j]), &bar_col[c], (int)n0,
                                (int)(x0 + c), (int)(oy * P.sy), (int)(cb * P.st_cblk));
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            /* S32 accumulator, U8 x U8, A MN-major (bit 15), B K-major, M = 128, N = npad */
            const uint32_t idesc = (2u << 4) | (1u << 15) | ((npad >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
            for (uint32_t cb = 0; cb < ncb; ++cb) {
                mbar_wait(&bar_w, cb & 1);
                uint32_t ready = 0; /* column slots known to have landed */
                for (uint32_t p = 0; p < S; ++p) {
                    for (uint32_t jx = 0; jx < fw; ++jx) {
                        while (ready <= p + jx) {
                            mbar_wait(&bar_col[ready], cb & 1);
                            ++ready;
                        }
                        asm volatile("tcgen05.fence::after_thread_sync;");
                        const uint32_t aC = smem_u32(sCol + (p + jx) * col_bytes);
                        const uint32_t aW = smem_u32(sW + jx * npad * 128u);
                        for (uint32_t k4 = 0; k4 < R / 32; ++k4) {
                            const uint64_t da = umma_desc(aC + k4 * 32 * TC_BN, 16, 1024); /* +32 k-rows of 128 plane bytes */
                            const uint64_t db = umma_desc(aW + k4 * 32, 16, 1024);         /* +32 B inside the swizzled weight row */
                            const uint32_t acc = (cb | jx | k4) ? 1u : 0u;
                            asm volatile(
                                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}" ::"r"(
                                    tmem_base + p * npad),
                                "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(0), "r"(0), "r"(0), "r"(0)
                                : "memory");
                        }
                    }
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar_free))
                             : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar_acc))
                         : "memory");
        }
    } else {
        /* epilogue, one output pixel after the other: TMEM lane = plane byte, column = pixel * npad + filter;
         * the reduced bytes are transposed through shared memory (the operand buffers are free once the
         * accumulators are complete) so that the global stores are 16 bytes per thread along a filter's row */
        mbar_wait(&bar_acc, 0);
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t sub = (uint32_t)warp & 3u;
        const uint32_t te = threadIdx.x - 64; /* 0..127 among the epilogue threads */
        const uint32_t l128 = sub * 32 + lane;
        const uint32_t n = n0 + l128;
        const uint32_t ncp = nc4 * 4;
        const uint32_t d4_first = n0 / (4 * P.Bp);
        uint8_t* s_out = smem;                                           /* [npad][128] */
        uint32_t* s_cst = reinterpret_cast<uint32_t*>(smem + 128 * 128); /* [npad][4] words of 4 digits */
        for (uint32_t i = te; i < npad * 4; i += 128) {
            const uint32_t f = i >> 2, d4 = d4_first + (i & 3u);
            s_cst[i] = d4 < nc4 ? __ldg(reinterpret_cast<const uint32_t*>(P.cst[j] + (uint64_t)f * ncp) + d4) : 0u;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const uint32_t q = (n < Nj ? n / (4 * P.Bp) : d4_first) - d4_first; /* < 4: a 128-byte chunk spans at most 3 digit words */
        const uint32_t sh = 8 * (n & 3u);
        const uint32_t col16 = (te & 7u) * 16;
        for (uint32_t p = 0; p < S; ++p) {
            const uint32_t pix = oy * P.ow + x0 + p;
            const bool live = x0 + p < P.ow && pix >= P.p0 && pix < P.p1; /* uniform over the CTA */
            if (!live) continue;
            for (uint32_t c0 = 0; c0 < npad; c0 += 32) {
                uint32_t v[32];
                const uint32_t taddr = tmem_base + ((sub * 32u) << 16) + p * npad + c0;
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                      "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                      "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const uint32_t f = c0 + i; /* filter */
                    uint32_t x = v[i] + ((s_cst[f * 4 + q] >> sh) & 0xffu);
                    x -= gc::mulhi32(x, m.mq) * m.q;
                    if ((int32_t)x < 0) x += m.q;
                    s_out[f * 128 + l128] = (uint8_t)x;
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            /* 8 threads x 16 bytes per filter row, 16 rows per pass */
            if (n0 + col16 < Nj) { /* Nj is a multiple of 64 */
                for (uint32_t f = te >> 3; f < npad; f += 16) {
                    if (f < P.v0 || f >= P.v1) continue;
                    const uint4 w = *reinterpret_cast<const uint4*>(s_out + f * 128 + col16);
                    *reinterpret_cast<uint4*>(P.out[j] + ((uint64_t)f * P.npix + pix) * Nj + n0 + col16) = w;
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory"); /* s_out is free for the next pixel */
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols));
}

/* ---- host side: tensor maps ---- */
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

static cudaError_t load_encode() {
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
        if (e != cudaSuccess) return e;
        if (!fn) return cudaErrorNotSupported;
        g_encode = (EncodeTiledFn)fn;
    }
    return cudaSuccess;
}

/* u8 row-major [outer][inner] (inner = contiguous bytes, multiple of 16), 128 x 128 boxes, zero fill out of bounds */
cudaError_t dense_tc_make_map(void* map_out, const void* ptr, uint64_t inner, uint64_t outer) {
    cudaError_t e = load_encode();
    if (e != cudaSuccess) return e;
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {inner};
    cuuint32_t box[2] = {128, 128};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode((CUtensorMap*)map_out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(ptr), dims, strides, box,
                          estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

/* input plane of a convolution: u8 [channel][row][column][row_bytes], boxes {128, f_w, f_h, cb} */
cudaError_t conv_tc_make_map(void* map_out, const void* ptr, uint64_t row_bytes, uint64_t in_w, uint64_t in_h,
                             uint64_t channels, uint32_t f_w, uint32_t f_h, uint32_t cb) {
    cudaError_t e = load_encode();
    if (e != cudaSuccess) return e;
    cuuint64_t dims[4] = {row_bytes, in_w, in_h, channels};
    cuuint64_t strides[3] = {row_bytes, row_bytes * in_w, row_bytes * in_w * in_h};
    cuuint32_t box[4] = {128, f_w, f_h, cb};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = g_encode((CUtensorMap*)map_out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(ptr), dims, strides, box,
                          estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

cudaError_t launch_dense_tc(const DenseTcMaps& maps, const DenseTcParams& p_in, cudaStream_t s) {
    DenseTcParams p = p_in;
    if (p.p1 == 0) p.p1 = p.npix ? p.npix : 1; /* no pixel range given: all of them */
    if (p.v1 <= p.v0 || p.p1 <= p.p0) return cudaSuccess;
    const uint32_t npix_l = p.p1 - p.p0;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_dense_tc<false, TC_STAGES_DENSE, TC_NB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)tc_smem_bytes(TC_STAGES_DENSE, TC_NB));
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_dense_tc<true, TC_STAGES_CONV, TC_NB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)tc_smem_bytes(TC_STAGES_CONV, TC_NB));
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    uint32_t max_cols = 0;
    for (int j = 0; j < p.k; ++j) max_cols = max(max_cols, ((p.mod[j].nc + 3) & ~3u) * p.Bp);
    const uint32_t mblocks = (p.v1 - p.v0 + TC_BM - 1) / TC_BM;
    if (p.conv && p.npad && p.strip) {
        static bool attr_s = false;
        const uint32_t smem_s = (p.strip + p.st_fw - 1) * p.st_rows * 128u + p.st_wboxes * 128u * 128u + 1024;
        if (!attr_s) {
            cudaError_t e = cudaFuncSetAttribute(k_conv_tc_strip, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e != cudaSuccess) return e;
            attr_s = true;
        }
        /* the output rows that hold pixels of [p0, p1), strips of `strip` pixels along each */
        const uint32_t rows = (p.p1 - 1) / p.ow - p.p0 / p.ow + 1, spr = (p.ow + p.strip - 1) / p.strip;
        dim3 grid((max_cols + TC_BN - 1) / TC_BN, rows * spr, p.k);
        k_conv_tc_strip<<<grid, TC_THREADS, smem_s, s>>>(maps, p);
        return cudaGetLastError();
    }
    if (p.conv && p.npad) {
        static bool attr_n = false;
        constexpr uint32_t smem_n = TC_STAGES_CONV * (TC_BK * TC_BN + TC_BM * TC_BK) + 1024;
        if (!attr_n) {
            cudaError_t e = cudaFuncSetAttribute(k_conv_tc_n<TC_STAGES_CONV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_n);
            if (e != cudaSuccess) return e;
            attr_n = true;
        }
        dim3 grid((max_cols + TC_BN - 1) / TC_BN, npix_l, p.k);
        k_conv_tc_n<TC_STAGES_CONV><<<grid, TC_THREADS, smem_n, s>>>(maps, p);
        return cudaGetLastError();
    }
    if (p.npix > 1 || p.conv) {
        dim3 grid((max_cols + TC_BN * TC_NB - 1) / (TC_BN * TC_NB), mblocks * npix_l, p.k);
        k_dense_tc<true, TC_STAGES_CONV, TC_NB><<<grid, TC_THREADS, tc_smem_bytes(TC_STAGES_CONV, TC_NB), s>>>(maps, p);
    } else {
        dim3 grid((max_cols + TC_BN * TC_NB - 1) / (TC_BN * TC_NB), mblocks, p.k);
        k_dense_tc<false, TC_STAGES_DENSE, TC_NB><<<grid, TC_THREADS, tc_smem_bytes(TC_STAGES_DENSE, TC_NB), s>>>(maps, p);
    }
    return cudaGetLastError();
}

}  // namespace dashb
