/*
 * kernels.cu -- hand-written sm_100a kernels of the B200 evaluator for
 * Dash's online garbled inference.  One thread evaluates one (element,
 * inference) pair of a gadget; the batch is the fastest-varying thread index
 * so that the 32 lanes of a warp share one element's ciphertext tables
 * (L1/L2-resident) and read/write 32 consecutive bytes of every digit row.
 *
 * Reference kernels / CPU routines each kernel replaces are cited at the
 * kernel (paths relative to dash/include/ of UzL-ITS/dash).
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <cstddef>
#include <cstring>

#include "../../include/dash_b200.h"
#include "kernels.h"

namespace dashb {

using gc::ModInfo;
using gc::u128;

/* ------------------------------------------------------------ AES tables */
/* Round keys of the fixed key 00..0f (crypto/cpu_aes_engine.h:23-25) as
 * big-endian words; uniform across a warp -> constant cache. */
__constant__ __align__(16) uint32_t c_rk[44];
/* the number 4, kept where the compiler cannot fold it (see te_at) */
__constant__ uint32_t c_four;
/* T-table Te0 and the S-box; copied into shared memory by every CTA that
 * hashes (cuda_aes_engine.h:27-32 keeps four tables in __constant__, which
 * serialises divergent lookups). */
__device__ uint32_t g_te0[256];
__device__ uint8_t g_sbox[256];

static uint32_t h_te[4][256];
static uint8_t h_sbox[256];
static uint32_t h_rk[44];
static bool h_tables_ready = false;

static void build_host_tables() {
    if (h_tables_ready) return;
    /* S-box from the multiplicative inverse in GF(2^8) + affine map (FIPS-197 5.1.1) */
    uint8_t p = 1, q = 1;
    do {
        p = (uint8_t)(p ^ (p << 1) ^ ((p & 0x80) ? 0x1b : 0));
        q ^= (uint8_t)(q << 1);
        q ^= (uint8_t)(q << 2);
        q ^= (uint8_t)(q << 4);
        if (q & 0x80) q ^= 0x09;
        uint8_t x = (uint8_t)(q ^ ((q << 1) | (q >> 7)) ^ ((q << 2) | (q >> 6)) ^ ((q << 3) | (q >> 5)) ^
                              ((q << 4) | (q >> 4)));
        h_sbox[p] = (uint8_t)(x ^ 0x63);
    } while (p != 1);
    h_sbox[0] = 0x63;
    for (int i = 0; i < 256; ++i) {
        uint32_t s = h_sbox[i];
        uint32_t s2 = ((s << 1) ^ ((s >> 7) * 0x1b)) & 0xff;
        uint32_t s3 = s2 ^ s;
        uint32_t t = (s2 << 24) | (s << 16) | (s << 8) | s3;
        h_te[0][i] = t;
        h_te[1][i] = (t >> 8) | (t << 24);
        h_te[2][i] = (t >> 16) | (t << 16);
        h_te[3][i] = (t >> 24) | (t << 8);
    }
    static const uint8_t rcon[10] = {0x01, 0x02, 0x04, 0x08, 0x10, 0x20, 0x40, 0x80, 0x1b, 0x36};
    for (int i = 0; i < 4; ++i)
        h_rk[i] = ((uint32_t)(4 * i) << 24) | ((uint32_t)(4 * i + 1) << 16) | ((uint32_t)(4 * i + 2) << 8) |
                  (uint32_t)(4 * i + 3);
    for (int i = 4; i < 44; ++i) {
        uint32_t t = h_rk[i - 1];
        if (i % 4 == 0) {
            t = (t << 8) | (t >> 24);
            t = ((uint32_t)h_sbox[t >> 24] << 24) | ((uint32_t)h_sbox[(t >> 16) & 0xff] << 16) |
                ((uint32_t)h_sbox[(t >> 8) & 0xff] << 8) | (uint32_t)h_sbox[t & 0xff];
            t ^= (uint32_t)rcon[i / 4 - 1] << 24;
        }
        h_rk[i] = h_rk[i - 4] ^ t;
    }
    h_tables_ready = true;
}

void host_aes_tables(const uint32_t** te0, const uint32_t** te1, const uint32_t** te2, const uint32_t** te3,
                     const uint8_t** sbox, const uint32_t** rk) {
    build_host_tables();
    *te0 = h_te[0];
    *te1 = h_te[1];
    *te2 = h_te[2];
    *te3 = h_te[3];
    *sbox = h_sbox;
    *rk = h_rk;
}

cudaError_t upload_aes_tables() {
    build_host_tables();
    cudaError_t e = cudaMemcpyToSymbol(c_rk, h_rk, sizeof(h_rk));
    if (e != cudaSuccess) return e;
    const uint32_t four = 4;
    e = cudaMemcpyToSymbol(c_four, &four, sizeof(four));
    if (e != cudaSuccess) return e;
    e = cudaMemcpyToSymbol(g_te0, h_te[0], sizeof(h_te[0]));
    if (e != cudaSuccess) return e;
    return cudaMemcpyToSymbol(g_sbox, h_sbox, sizeof(h_sbox));
}

/* Shared-memory copy of the hash tables of one CTA. */
struct AesSmem {
    uint32_t te0[256], te1[256], te2[256], te3[256];
    uint8_t sbox[256];
};

__device__ __forceinline__ void aes_smem_load(AesSmem& a) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        uint32_t t = g_te0[i];
        a.te0[i] = t;
        a.te1[i] = (t >> 8) | (t << 24);
        a.te2[i] = (t >> 16) | (t << 16);
        a.te3[i] = (t >> 24) | (t << 8);
        a.sbox[i] = g_sbox[i];
    }
    __syncthreads();
}

/* Copies a ModInfo out of the (dynamically indexed) kernel-parameter bank
 * into registers and hides its origin from the compiler, which would
 * otherwise re-read the constant bank (4 instructions) at every use. */
__device__ __forceinline__ ModInfo pin(const ModInfo& src) {
    ModInfo m = src;
    asm volatile("" : "+r"(m.q), "+r"(m.nc), "+r"(m.t), "+r"(m.nchunks), "+r"(m.Q), "+r"(m.mq));
    asm volatile("" : "+r"(m.lim), "+r"(m.top_len), "+r"(m.s), "+r"(m.dn), "+r"(m.v), "+r"(m.limbs), "+r"(m.pow2s));
    asm volatile("" : "+r"(m.cw), "+r"(m.cwf), "+r"(m.q4), "+r"(m.Qw));
    return m;
}

/* Te<TAB>[byte I of s].  The byte comes out with one PRMT (ALU pipe); the
 * scaling to a table offset is a multiply by a constant the compiler cannot
 * see (c_four, constant bank) so that it stays an IMAD on the FMA pipe: the
 * ALU pipe is the busy one in the gadget kernels (ncu), and the usual
 * shift+mask form puts two ALU operations on every one of the 160 lookups. */
template <int TAB, int I>
__device__ __forceinline__ uint32_t te_at(const AesSmem& a, uint32_t s, uint32_t four) {
    const uint32_t byte = I == 0 ? (s & 0xffu) : __byte_perm(s, 0, 0x4440 + I);
    const char* tab = reinterpret_cast<const char*>(TAB == 0 ? a.te0 : TAB == 1 ? a.te1 : TAB == 2 ? a.te2 : a.te3);
    return *reinterpret_cast<const uint32_t*>(tab + byte * four);
}

template <int I>
__device__ __forceinline__ uint32_t sbox_at(const AesSmem& a, uint32_t s) {
    return a.sbox[I == 0 ? (s & 0xffu) : __byte_perm(s, 0, 0x4440 + I)];
}

/* H(x): AES-128 under the fixed key of x as 16 little-endian bytes
 * (crypto/cpu_aes_engine.h:23-40).  Same function as gc::aes128_encrypt,
 * which stays the host-checkable statement of the cipher.  One copy per
 * kernel (not inlined at its ~8 call sites) with rolled rounds: the gadget
 * kernels then fit the instruction cache (ncu: 24 % stall_no_inst when
 * inlined and unrolled). */
#ifndef DASHB_AES_UNROLL
#define DASHB_AES_UNROLL 1
#endif
constexpr int AES_UNROLL = DASHB_AES_UNROLL; /* rounds per trip of the middle-round loop (1, 3 or 9) */
#ifndef DASHB_HASH_INLINE
#define DASHB_HASH_INLINE 0
#endif
#if DASHB_HASH_INLINE
__device__ __forceinline__
#else
__device__ __noinline__
#endif
u128 hash(const AesSmem& a, u128 x) {
    const uint32_t four = c_four;
    uint32_t s0 = gc::bswap32((uint32_t)x.lo) ^ c_rk[0];
    uint32_t s1 = gc::bswap32((uint32_t)(x.lo >> 32)) ^ c_rk[1];
    uint32_t s2 = gc::bswap32((uint32_t)x.hi) ^ c_rk[2];
    uint32_t s3 = gc::bswap32((uint32_t)(x.hi >> 32)) ^ c_rk[3];
#pragma unroll(AES_UNROLL)
    for (int r = 1; r < 10; ++r) {
        const uint4 k4 = reinterpret_cast<const uint4*>(c_rk)[r];
        const uint32_t t0 = te_at<0, 3>(a, s0, four) ^ te_at<1, 2>(a, s1, four) ^ te_at<2, 1>(a, s2, four) ^ te_at<3, 0>(a, s3, four) ^ k4.x;
        const uint32_t t1 = te_at<0, 3>(a, s1, four) ^ te_at<1, 2>(a, s2, four) ^ te_at<2, 1>(a, s3, four) ^ te_at<3, 0>(a, s0, four) ^ k4.y;
        const uint32_t t2 = te_at<0, 3>(a, s2, four) ^ te_at<1, 2>(a, s3, four) ^ te_at<2, 1>(a, s0, four) ^ te_at<3, 0>(a, s1, four) ^ k4.z;
        const uint32_t t3 = te_at<0, 3>(a, s3, four) ^ te_at<1, 2>(a, s0, four) ^ te_at<2, 1>(a, s1, four) ^ te_at<3, 0>(a, s2, four) ^ k4.w;
        s0 = t0, s1 = t1, s2 = t2, s3 = t3;
    }
    const uint32_t o0 = (sbox_at<3>(a, s0) << 24) ^ (sbox_at<2>(a, s1) << 16) ^ (sbox_at<1>(a, s2) << 8) ^ sbox_at<0>(a, s3) ^ c_rk[40];
    const uint32_t o1 = (sbox_at<3>(a, s1) << 24) ^ (sbox_at<2>(a, s2) << 16) ^ (sbox_at<1>(a, s3) << 8) ^ sbox_at<0>(a, s0) ^ c_rk[41];
    const uint32_t o2 = (sbox_at<3>(a, s2) << 24) ^ (sbox_at<2>(a, s3) << 16) ^ (sbox_at<1>(a, s0) << 8) ^ sbox_at<0>(a, s1) ^ c_rk[42];
    const uint32_t o3 = (sbox_at<3>(a, s3) << 24) ^ (sbox_at<2>(a, s0) << 16) ^ (sbox_at<1>(a, s1) << 8) ^ sbox_at<0>(a, s2) ^ c_rk[43];
    u128 out;
    out.lo = (uint64_t)gc::bswap32(o0) | ((uint64_t)gc::bswap32(o1) << 32);
    out.hi = (uint64_t)gc::bswap32(o2) | ((uint64_t)gc::bswap32(o3) << 32);
    return out;
}

__device__ __forceinline__ u128 ldg128(const u128* p) {
    uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    u128 r;
    r.lo = (uint64_t)v.x | ((uint64_t)v.y << 32);
    r.hi = (uint64_t)v.z | ((uint64_t)v.w << 32);
    return r;
}

/* Plane addressing (see kernels.h): digits are stored four to a 32-bit
 * word, word (e, d/4, b) at e*row + (d/4)*4*Bp + 4*b, so a thread reads or
 * writes 4 digits per access and a warp 128 contiguous bytes. */
__device__ __forceinline__ uint32_t nc4_of(const ModInfo& m) { return (m.nc + 3) >> 2; }
__device__ __forceinline__ uint64_t plane_off(const ModInfo& m, uint64_t e, uint32_t b, uint32_t Bp) {
    return (e * nc4_of(m) * (uint64_t)Bp + b) * 4;
}

/* compress the label (element e, inference b) of a digit plane.  The words
 * are read two ahead of the Horner steps that consume them: a rolled loop
 * that used each word right after its load exposed the full memory latency
 * once per word (ncu: the largest long-scoreboard stall after the half gates'
 * re-reads).  A word's four digits enter the value at once
 * (gc::WordCompressor: two dp4a + two multiply-adds per word instead of four
 * Horner steps with their chunk bookkeeping). */
__device__ __forceinline__ u128 load_compress(const uint8_t* plane, const ModInfo& m, uint64_t e, uint32_t b,
                                               uint32_t Bp, uint32_t& colour) {
    const uint8_t* p = plane + plane_off(m, e, b, Bp);
    const int top = (int)nc4_of(m) - 1;
    auto word = [&](int d4) { return d4 >= 0 ? *reinterpret_cast<const uint32_t*>(p + (uint64_t)d4 * 4 * Bp) : 0u; };
    uint32_t w = 0, n0 = word(top), n1 = word(top - 1);
    if (m.cw) {
        /* whatever a producer left in the padding bytes of the top word stays out */
        n0 &= 0xffffffffu >> (8 * ((4u - (m.nc & 3u)) & 3u));
        gc::WordCompressor c(m);
#pragma unroll 1
        for (int d4 = top; d4 >= 0; --d4) {
            w = n0;
            n0 = n1;
            n1 = word(d4 - 2);
            c.push(w);
        }
        colour = w & 0xff;
        return c.value();
    }
    gc::Compressor c(m);
    uint32_t valid = m.nc - 4 * (uint32_t)top; /* digits in the top word */
#pragma unroll 1
    for (int d4 = top; d4 >= 0; --d4) {
        w = n0;
        n0 = n1;
        n1 = word(d4 - 2);
        if (valid > 3) c.push(w >> 24);
        if (valid > 2) c.push((w >> 16) & 0xff);
        if (valid > 1) c.push((w >> 8) & 0xff);
        c.push(w & 0xff);
        valid = 4;
    }
    colour = w & 0xff;
    return c.value();
}

/* write the digits of compressed label x into a plane */
__device__ __forceinline__ void store_digits(uint8_t* plane, const ModInfo& m, uint64_t e, uint32_t b, uint32_t Bp,
                                             u128 x) {
    uint8_t* p = plane + plane_off(m, e, b, Bp);
    uint32_t w = 0;
    const uint32_t last = m.nc - 1;
    gc::for_each_digit(x, m, [&](uint32_t d, uint32_t v) {
        w |= v << (8 * (d & 3));
        if ((d & 3) == 3 || d == last) {
            *reinterpret_cast<uint32_t*>(p + (uint64_t)(d >> 2) * 4 * Bp) = w;
            w = 0;
        }
    });
}

/* --------------------------------------------------- ingest / egress */
/* Host layout int16 [batch][label][nc] (LabelTensor order,
 * garbling/label_tensor.h:238-246) <-> device plane.  One CTA moves one
 * label of 32 inferences through a shared-memory tile so that both sides are
 * coalesced.  Replaces the per-modulus cudaMemcpy of cuda_move_inputs /
 * cuda_move_outputs (garbled_circuit_interface.h:551-565,733-736). */
__global__ void __launch_bounds__(128) k_ingest(const int16_t* __restrict__ src, uint8_t* __restrict__ plane,
                                                uint64_t n, uint32_t nc, uint32_t batch, uint32_t Bp) {
    __shared__ uint32_t tile[32][33]; /* [b][d4] */
    const uint64_t e = blockIdx.x;
    const uint32_t b0 = blockIdx.y * 32;
    const uint32_t nc4 = (nc + 3) >> 2;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t i = warp; i < 32; i += 4) {
        const uint32_t b = b0 + i;
        const int16_t* row = src + ((uint64_t)b * n + e) * nc;
        for (uint32_t d4 = lane; d4 < nc4; d4 += 32) {
            uint32_t w = 0;
            if (b < batch)
                for (uint32_t t = 0; t < 4 && 4 * d4 + t < nc; ++t) w |= (uint32_t)(uint8_t)row[4 * d4 + t] << (8 * t);
            tile[i][d4] = w;
        }
    }
    __syncthreads();
    for (uint32_t d4 = warp; d4 < nc4; d4 += 4) {
        const uint32_t b = b0 + lane;
        if (b < Bp) *reinterpret_cast<uint32_t*>(plane + ((e * nc4 + d4) * (uint64_t)Bp + b) * 4) = tile[lane][d4];
    }
}

__global__ void __launch_bounds__(128) k_egress(const uint8_t* __restrict__ plane, int16_t* __restrict__ dst,
                                                uint64_t n, uint32_t nc, uint32_t batch, uint32_t Bp) {
    __shared__ uint32_t tile[32][33];
    const uint64_t e = blockIdx.x;
    const uint32_t b0 = blockIdx.y * 32;
    const uint32_t nc4 = (nc + 3) >> 2;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t d4 = warp; d4 < nc4; d4 += 4) {
        const uint32_t b = b0 + lane;
        tile[lane][d4] = b < Bp ? *reinterpret_cast<const uint32_t*>(plane + ((e * nc4 + d4) * (uint64_t)Bp + b) * 4) : 0;
    }
    __syncthreads();
    for (uint32_t i = warp; i < 32; i += 4) {
        const uint32_t b = b0 + i;
        if (b >= batch) continue;
        int16_t* row = dst + ((uint64_t)b * n + e) * nc;
        for (uint32_t d = lane; d < nc; d += 32) row[d] = (int16_t)((tile[i][d >> 2] >> (8 * (d & 3))) & 0xff);
    }
}

/* k_egress for all moduli at once (blockIdx.z): one launch per lane instead of k -- at the end of a step every
 * launch queues behind the other lanes' grids -- and a destination that may be the caller's mapped host buffer
 * (2-byte stores, 64 contiguous bytes per warp; the outputs of a step are a few MB). */
__global__ void __launch_bounds__(128) k_egress_all(const __grid_constant__ EgressParams P) {
    __shared__ uint32_t tile[32][33];
    const int j = blockIdx.z;
    const uint32_t nc = P.nc[j], nc4 = (nc + 3) >> 2;
    const uint8_t* plane = P.plane[j];
    const uint64_t e = blockIdx.x;
    const uint32_t b0 = blockIdx.y * 32;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t d4 = warp; d4 < nc4; d4 += 4) {
        const uint32_t b = b0 + lane;
        tile[lane][d4] = b < P.Bp ? *reinterpret_cast<const uint32_t*>(plane + ((e * nc4 + d4) * (uint64_t)P.Bp + b) * 4) : 0;
    }
    __syncthreads();
    for (uint32_t i = warp; i < 32; i += 4) {
        const uint32_t b = b0 + i;
        if (b >= P.batch) continue;
        int16_t* row = P.dst[j] + ((uint64_t)b * P.n + e) * nc;
        for (uint32_t d = lane; d < nc; d += 32) row[d] = (int16_t)((tile[i][d >> 2] >> (8 * (d & 3))) & 0xff);
    }
}

cudaError_t launch_egress_all(const EgressParams& p, cudaStream_t s) {
    if (p.n == 0 || p.batch == 0) return cudaSuccess;
    dim3 grid((unsigned)p.n, (p.Bp + 31) / 32, (unsigned)p.k);
    if (p.n > 0x7fffffffull || grid.y > 65535u) return cudaErrorInvalidValue;
    k_egress_all<<<grid, 128, 0, s>>>(p);
    return cudaGetLastError();
}

/*
 * The same layout change without the staging copy: the caller's pinned host
 * buffers are mapped into the device's address space, so the kernel reads
 * them over PCIe itself (cuda_move_inputs' cudaMemcpy and the k_ingest pass
 * in one).  A CTA takes one inference's labels of ING_E consecutive elements
 * of one modulus -- one contiguous span of the host buffer -- with 16-byte
 * loads, every host byte exactly once, through shared memory; the words of
 * four u8 digits then go to the plane (4-byte stores, one sector each: they
 * merge in L2 with the other inferences' stores of the same rows).  PCIe is
 * the bound: the copy engine needs ~190 MB per copy to reach the link rate
 * (profiles/r1_v9_summary.md), loads issued from a few dozen CTAs do not.
 */
constexpr int ING_E = 32;
/* 128 threads at <= 32 registers: such a CTA fits on an SM next to eight
 * k_sign CTAs (8 x 128 x 60 registers) instead of displacing one of them */
constexpr int ING_THREADS = 128;
__global__ void __launch_bounds__(ING_THREADS, 16) k_ingest_direct(const __grid_constant__ IngestParams P) {
    __shared__ uint4 buf[(ING_E * 128 * 2 + 32) / 16 + 1];
    /* A small persistent grid: the kernel shares the GPU with the lanes that
     * already compute and only has to keep the PCIe link busy (a few hundred
     * KB in flight), not to fill the machine -- one CTA per work item took the
     * SM slots away from the gadget kernels (measured: step 5.6 -> 8.1 ms). */
    const uint32_t nchunk = (uint32_t)((P.n + ING_E - 1) / ING_E);
    const uint64_t items = (uint64_t)nchunk * P.batch * P.k;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint64_t item = blockIdx.x; item < items; item += gridDim.x) {
        /* modulus slowest, then inference, then element chunk: a CTA's items walk the host buffer forward */
        const int j = (int)(item / ((uint64_t)nchunk * P.batch));
        const uint64_t rem = item - (uint64_t)j * nchunk * P.batch;
        const uint32_t b = (uint32_t)(rem / nchunk);
        const uint64_t e0 = (rem - (uint64_t)b * nchunk) * ING_E;
        const uint32_t ne = (uint32_t)min((uint64_t)ING_E, P.n - e0);
        const uint32_t nc = P.nc[j], nc4 = (nc + 3) >> 2;
        const char* src = reinterpret_cast<const char*>(P.src[j] + ((uint64_t)b * P.n + e0) * nc);
        const char* end = reinterpret_cast<const char*>(P.end[j]);
        const uint32_t bytes = ne * nc * 2;
        const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 15u);
        const char* start = src - mis; /* 16-byte aligned; at worst 14 bytes before the span, in the same page */
        const uint32_t nvec = (bytes + mis + 15) >> 4;
        for (uint32_t i = threadIdx.x; i < nvec; i += ING_THREADS) {
            const char* p = start + 16ull * i;
            uint4 v;
            if (p + 16 <= end) {
                v = __ldcs(reinterpret_cast<const uint4*>(p)); /* streaming: read once */
            } else { /* the very last vector of the caller's buffer: do not read past its end */
                uint16_t h[8];
                for (int t = 0; t < 8; ++t) h[t] = (p + 2 * t + 2 <= end) ? *reinterpret_cast<const uint16_t*>(p + 2 * t) : (uint16_t)0;
                v.x = h[0] | ((uint32_t)h[1] << 16);
                v.y = h[2] | ((uint32_t)h[3] << 16);
                v.z = h[4] | ((uint32_t)h[5] << 16);
                v.w = h[6] | ((uint32_t)h[7] << 16);
            }
            buf[i] = v;
        }
        __syncthreads();
        const uint16_t* dig = reinterpret_cast<const uint16_t*>(reinterpret_cast<const char*>(buf) + mis);
        uint8_t* plane = P.plane[j];
        /* the inference columns past the batch (up to Bp) are kept zero by the items of the last inference */
        const uint32_t nb = (b + 1 == P.batch) ? P.Bp - b : 1;
        /* a warp walks the words of one element at a time */
        for (uint32_t el = warp; el < ne; el += ING_THREADS / 32) {
            const uint16_t* row = dig + el * nc;
            for (uint32_t d4 = lane; d4 < nc4; d4 += 32) {
                uint32_t w = 0;
#pragma unroll
                for (uint32_t t = 0; t < 4; ++t)
                    if (4 * d4 + t < nc) w |= (uint32_t)(uint8_t)row[4 * d4 + t] << (8 * t);
                uint32_t* dst = reinterpret_cast<uint32_t*>(plane + (((e0 + el) * nc4 + d4) * (uint64_t)P.Bp + b) * 4);
                __stcs(dst, w); /* streaming: the planes must not push the ciphertext tables out of L2 */
                for (uint32_t x = 1; x < nb; ++x) __stcs(dst + x, 0u);
            }
        }
        __syncthreads(); /* the buffer is free for the next item */
    }
}

cudaError_t launch_ingest_direct(const IngestParams& p, cudaStream_t s) {
    if (p.n == 0 || p.batch == 0) return cudaSuccess;
    static const int ctas = getenv("DASHB_INGEST_CTAS") ? atoi(getenv("DASHB_INGEST_CTAS")) : 192;
    const uint64_t items = ((p.n + ING_E - 1) / ING_E) * p.batch * p.k;
    const unsigned grid = (unsigned)std::min<uint64_t>(items, (uint64_t)std::max(1, ctas));
    k_ingest_direct<<<grid, ING_THREADS, 0, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_ingest(const int16_t* src, uint8_t* plane, uint64_t n, uint32_t nc, uint32_t batch, uint32_t Bp,
                          cudaStream_t s) {
    dim3 grid((unsigned)n, (Bp + 31) / 32);
    k_ingest<<<grid, 128, 0, s>>>(src, plane, n, nc, batch, Bp);
    return cudaGetLastError();
}
cudaError_t launch_egress(const uint8_t* plane, int16_t* dst, uint64_t n, uint32_t nc, uint32_t batch, uint32_t Bp,
                          cudaStream_t s) {
    dim3 grid((unsigned)n, (Bp + 31) / 32);
    k_egress<<<grid, 128, 0, s>>>(plane, dst, n, nc, batch, Bp);
    return cudaGetLastError();
}

/* ------------------------------------------------------------- linear */
/*
 * Garbled dense / conv2d for all CRT moduli in one launch:
 *   out[v*npix+pix][d][b] = ( sum_k Wq[v][k] * in[base(pix)+koff[k]][d][b]
 *                             + cst[v][d] ) mod q
 * with Wq = W mod q and cst = bias label + (#zero residues in row v) * zero
 * label -- the exact reformulation of LabelTensor::matvecmul[_tf] /
 * conv2d with their zero-weight rule (garbling/label_tensor.h:783-858,
 * 920-1002) and of MatMulModAddZero[TF] / Conv2dModZero
 * (misc/cuda_util.h:95-138,209-268).  u8 x u8 -> u32 dot products with
 * dp4a: a thread owns 4 adjacent columns (one 32-bit word of a digit row)
 * and 8 output rows; four input rows are byte-transposed in registers into
 * K-packed words.  K <= 65536 keeps the u32 accumulators exact.
 */
constexpr int LIN_THREADS = 128;
constexpr int LIN_ROWS = 8;
constexpr int LIN_KC = 1024; /* taps staged in shared memory per chunk */

constexpr int LIN_PIXC = 4; /* output pixels a CTA walks through with one staging of the weights */

/* The multiply-accumulate loop over one staged chunk of taps, for a row
 * group of which only the first NR rows exist (5 filters: 20 dot products per
 * tap group instead of 32). */
template <int NR>
__device__ __forceinline__ void lin_taps(const uint8_t* in, uint64_t rowlen, const int32_t* s_koff, const uint4 (*s_w)[2],
                                         uint32_t k4_n, uint32_t (&acc)[LIN_ROWS][4]) {
#pragma unroll 1
    for (uint32_t kk = 0; kk < k4_n; ++kk) {
        const uint32_t x0 = __ldg(reinterpret_cast<const uint32_t*>(in + (int64_t)s_koff[4 * kk + 0] * rowlen));
        const uint32_t x1 = __ldg(reinterpret_cast<const uint32_t*>(in + (int64_t)s_koff[4 * kk + 1] * rowlen));
        const uint32_t x2 = __ldg(reinterpret_cast<const uint32_t*>(in + (int64_t)s_koff[4 * kk + 2] * rowlen));
        const uint32_t x3 = __ldg(reinterpret_cast<const uint32_t*>(in + (int64_t)s_koff[4 * kk + 3] * rowlen));
        /* 4x4 byte transpose: y_c = (x0.c, x1.c, x2.c, x3.c) */
        const uint32_t t0 = __byte_perm(x0, x1, 0x5140), t1 = __byte_perm(x0, x1, 0x7362);
        const uint32_t u0 = __byte_perm(x2, x3, 0x5140), u1 = __byte_perm(x2, x3, 0x7362);
        const uint32_t y0 = __byte_perm(t0, u0, 0x5410), y1 = __byte_perm(t0, u0, 0x7632);
        const uint32_t y2 = __byte_perm(t1, u1, 0x5410), y3 = __byte_perm(t1, u1, 0x7632);
        const uint4 wa = s_w[kk][0];
        uint4 wb = make_uint4(0, 0, 0, 0);
        if (NR > 4) wb = s_w[kk][1];
        const uint32_t w[LIN_ROWS] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            acc[r][0] = __dp4a(y0, w[r], acc[r][0]);
            acc[r][1] = __dp4a(y1, w[r], acc[r][1]);
            acc[r][2] = __dp4a(y2, w[r], acc[r][2]);
            acc[r][3] = __dp4a(y3, w[r], acc[r][3]);
        }
    }
}

__global__ void __launch_bounds__(LIN_THREADS, 8) k_linear(const __grid_constant__ LinearParams P) {
    __shared__ uint4 s_w[LIN_KC / 4][2];  /* [k/4][8 rows] packed weights */
    __shared__ int32_t s_koff[LIN_KC];
    const int j = blockIdx.z;
    const ModInfo m = pin(P.mod[j]);
    const uint32_t ncols = ((m.nc + 3) & ~3u) * P.Bp; /* nc4 words of 4 digits per inference */
    const uint32_t col = (blockIdx.x * LIN_THREADS + threadIdx.x) * 4;
    /* the grid's x extent is sized for the modulus with the most digits: CTAs
     * past this modulus' columns leave before staging anything (58 % of the
     * CTAs of a k = 9 launch) */
    if (blockIdx.x * LIN_THREADS * 4 >= ncols) return;
    /* groups of 8 rows that overlap [v0, v1) */
    const uint32_t vg_first = P.v0 / LIN_ROWS;
    const uint32_t nvg = (P.v1 + LIN_ROWS - 1) / LIN_ROWS - vg_first;
    const uint32_t by = P.y0 + blockIdx.y;
    const uint32_t vg = vg_first + by % nvg;
    const uint32_t nr = min((uint32_t)LIN_ROWS, P.v1 - vg * LIN_ROWS); /* rows of this group that exist */
    const uint32_t pix0 = P.p0 + (by / nvg) * LIN_PIXC;
    const uint32_t pix1 = min(pix0 + LIN_PIXC, P.p1);
    const uint64_t rowlen = ncols;
    const bool active = col < ncols;
    const uint32_t k4_total = P.Kp / 4;
    const bool one_chunk = k4_total <= LIN_KC / 4;
    /* the thread's 4 columns are digits 4*d4 .. 4*d4+3 of one inference */
    const uint32_t d4 = col / (4 * P.Bp);
    const uint32_t nc4 = (m.nc + 3) >> 2;

    auto stage = [&](uint32_t k4_0, uint32_t k4_n) {
        const uint4* wsrc = reinterpret_cast<const uint4*>(P.w4[j]) + ((uint64_t)vg * k4_total + k4_0) * 2;
        for (uint32_t i = threadIdx.x; i < k4_n * 2; i += LIN_THREADS) (&s_w[0][0])[i] = __ldg(wsrc + i);
        for (uint32_t i = threadIdx.x; i < k4_n * 4; i += LIN_THREADS) s_koff[i] = __ldg(P.koff + k4_0 * 4 + i);
    };
    if (one_chunk) {
        stage(0, k4_total);
        __syncthreads();
    }
    for (uint32_t pix = pix0; pix < pix1; ++pix) {
        const uint64_t base = (uint64_t)(pix / P.ow) * P.step_y + (uint64_t)(pix % P.ow) * P.step_x;
        const uint8_t* in = P.in[j] + base * rowlen + col;
        uint32_t acc[LIN_ROWS][4];
#pragma unroll
        for (int r = 0; r < LIN_ROWS; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[r][c] = 0;

        for (uint32_t k4_0 = 0; k4_0 < k4_total; k4_0 += LIN_KC / 4) {
            const uint32_t k4_n = min((uint32_t)(LIN_KC / 4), k4_total - k4_0);
            if (!one_chunk) {
                __syncthreads();
                stage(k4_0, k4_n);
                __syncthreads();
            }
            if (!active) continue;
            if (nr > 5)
                lin_taps<8>(in, rowlen, s_koff, s_w, k4_n, acc);
            else if (nr > 2)
                lin_taps<5>(in, rowlen, s_koff, s_w, k4_n, acc);
            else
                lin_taps<2>(in, rowlen, s_koff, s_w, k4_n, acc);
        }
        if (!active) continue;
        /* row v of this pixel: output word and constant word step by a fixed
         * stride from row to row */
        const uint64_t ostep = (uint64_t)P.npix * rowlen;
        uint8_t* op = P.out[j] + ((uint64_t)(vg * LIN_ROWS) * P.npix + pix) * rowlen + col;
        const uint32_t* cp = reinterpret_cast<const uint32_t*>(P.cst[j]) + (uint64_t)(vg * LIN_ROWS) * nc4 + d4;
#pragma unroll
        for (int r = 0; r < LIN_ROWS; ++r, op += ostep, cp += nc4) {
            const uint32_t v = vg * LIN_ROWS + r;
            if (v >= P.v1) break;
            if (v < P.v0) continue;
            const uint32_t cst = __ldg(cp);
            uint32_t o = 0;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t x = acc[r][c] + ((cst >> (8 * c)) & 0xff);
                x -= gc::mulhi32(x, m.mq) * m.q;
                /* x mod q for any 32-bit x: the reciprocal over-estimates the
                 * quotient by at most one; never when the launch's largest
                 * possible sum is within the reciprocal's exact range */
                if (!P.exact_mod && (int32_t)x < 0) x += m.q;
                o |= x << (8 * c);
            }
            *reinterpret_cast<uint32_t*>(op) = o;
        }
    }
}

cudaError_t launch_linear(const LinearParams& p, cudaStream_t s) {
    uint32_t max_cols = 0;
    for (int j = 0; j < p.k; ++j) max_cols = max(max_cols, ((p.mod[j].nc + 3) & ~3u) * p.Bp);
    if (p.v1 <= p.v0) return cudaSuccess; /* this rank owns no row of the layer */
    const uint32_t nvg = (p.v1 + LIN_ROWS - 1) / LIN_ROWS - p.v0 / LIN_ROWS;
    LinearParams q = p;
    q.exact_mod = 1; /* K (q-1)^2 + (q-1) <= lim for every modulus: no fix-up after the reciprocal */
    for (int j = 0; j < p.k; ++j) {
        const uint64_t qm = p.mod[j].q - 1;
        if ((uint64_t)p.K * qm * qm + qm > p.mod[j].lim) q.exact_mod = 0;
    }
    /* (pixel group, row group) pairs sit on grid.y, which ends at 65535: a large
     * convolution (222x222 outputs, 64 filters: 98 568 pairs) goes out as several launches */
    if (q.p1 == 0) q.p1 = p.npix; /* no pixel range given: all of them */
    if (q.p1 <= q.p0) return cudaSuccess;
    const uint64_t ny = (uint64_t)((q.p1 - q.p0 + LIN_PIXC - 1) / LIN_PIXC) * nvg;
    for (uint64_t y0 = 0; y0 < ny; y0 += 65535) {
        q.y0 = (uint32_t)y0;
        dim3 grid((max_cols / 4 + LIN_THREADS - 1) / LIN_THREADS, (unsigned)std::min<uint64_t>(65535, ny - y0), p.k);
        k_linear<<<grid, LIN_THREADS, 0, s>>>(q);
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

/* Compressed wire form: a label crosses PCIe as the 16
 * bytes LabelTensor::compress produces (garbling/label_tensor.h:715-727)
 * instead of 2*nc bytes of int16 digits; host layout [batch][label]. */
__global__ void __launch_bounds__(128) k_ingest_compressed(const u128* __restrict__ src, uint8_t* __restrict__ plane, uint64_t n,
                                                            const ModInfo m, uint32_t batch, uint32_t Bp) {
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * batch) return;
    const uint32_t b = (uint32_t)(idx % batch);
    const uint64_t e = idx / batch;
    store_digits(plane, m, e, b, Bp, ldg128(src + (uint64_t)b * n + e));
}

__global__ void __launch_bounds__(128) k_egress_compressed(const uint8_t* __restrict__ plane, u128* __restrict__ dst, uint64_t n,
                                                            const ModInfo m, uint32_t batch, uint32_t Bp) {
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * batch) return;
    const uint32_t b = (uint32_t)(idx % batch);
    const uint64_t e = idx / batch;
    uint32_t colour;
    dst[(uint64_t)b * n + e] = load_compress(plane, m, e, b, Bp, colour);
}

cudaError_t launch_ingest_compressed(const void* src, uint8_t* plane, uint64_t n, const ModInfo& m, uint32_t batch, uint32_t Bp,
                                     cudaStream_t s) {
    if (n * batch == 0) return cudaSuccess;
    k_ingest_compressed<<<(unsigned)((n * batch + 127) / 128), 128, 0, s>>>((const u128*)src, plane, n, m, batch, Bp);
    return cudaGetLastError();
}

cudaError_t launch_egress_compressed(const uint8_t* plane, void* dst, uint64_t n, const ModInfo& m, uint32_t batch, uint32_t Bp,
                                     cudaStream_t s) {
    if (n * batch == 0) return cudaSuccess;
    k_egress_compressed<<<(unsigned)((n * batch + 127) / 128), 128, 0, s>>>(plane, (u128*)dst, n, m, batch, Bp);
    return cudaGetLastError();
}

/* -------------------------------------------------------- projection */
/* GarbledProjection::cuda_evaluate / eval_proj (layer/garbled_projection.h:
 * 168-194, misc/cuda_util.h:297-317): out = digits(ciphers[e][colour] -
 * H(compress(in))). */
__global__ void __launch_bounds__(128) k_projection(const __grid_constant__ ProjParams P) {
    __shared__ AesSmem aes;
    aes_smem_load(aes);
    const int j = blockIdx.y;
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.n * P.batch) return;
    const uint32_t b = (uint32_t)(idx % P.batch);
    const uint64_t e = slice_elem(P.e0, P.seg, P.stride, idx / P.batch);
    uint32_t colour;
    const ModInfo mi = pin(P.min[j]), mo = pin(P.mout[j]);
    const u128 key = load_compress(P.in[j], mi, e, b, P.Bp, colour);
    const u128 h = hash(aes, key);
    const u128 row = ldg128(P.ciphers[j] + e * mi.q + colour);
    store_digits(P.out[j], mo, e, b, P.Bp, gc::sub128(row, h));
}

cudaError_t launch_projection(const ProjParams& p, cudaStream_t s) {
    if (p.n == 0) return cudaSuccess;
    const uint64_t total = p.n * p.batch;
    dim3 grid((unsigned)((total + 127) / 128), p.k);
    k_projection<<<grid, 128, 0, s>>>(p);
    return cudaGetLastError();
}

/* ---------------------------------------- projection, allocation-free */
/*
 * The digit planes of dashb_projection_static.  This is the only device
 * memory that entry point uses besides the caller's own buffers; it is part
 * of the library image, so every GPU that loads the library carries these
 * 8 MiB until the process ends.  In host code the name below is only a
 * stand-in for the symbol: the address kernels may dereference comes from
 * cudaGetSymbolAddress (static_scratch_address).
 */
static __device__ __align__(256) unsigned char s_static_scratch[DASHB_STATIC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
static unsigned char* g_static_scratch_addr[MAX_DEVICES]; /* device address per device, 0 = not looked up yet */

/* One-off per device, with that device current. */
static cudaError_t static_scratch_address(int device, unsigned char** out) {
    if (!g_static_scratch_addr[device]) {
        unsigned char* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&ptr, s_static_scratch);
        if (e != cudaSuccess) return e;
        g_static_scratch_addr[device] = ptr;
    }
    *out = g_static_scratch_addr[device];
    return cudaSuccess;
}

}  // namespace dashb

/* See include/dash_b200.h.  Same three kernels as the allocated path
 * (k_ingest -> k_projection -> k_egress); only the planes in between live in
 * s_static_scratch instead of cudaMalloc'ed buffers. */
extern "C" int dashb_projection_static(const int16_t* d_in, const dashb_u128* d_ciphers, int16_t* d_out, uint64_t n,
                                       uint64_t batch, int q_in, int q_out, void* stream) {
    using namespace dashb;
    if (!d_in || !d_ciphers || !d_out) return report_error(DASHB_EINVAL, "dashb_projection_static: null device pointer");
    if (q_in < 2 || q_in > 256 || q_out < 2 || q_out > 256)
        return report_error(DASHB_EINVAL, "dashb_projection_static: moduli must lie in [2,256] (got %d -> %d)", q_in, q_out);
    if (n == 0 || batch == 0 || batch > (1u << 24) || n > (1u << 30))
        return report_error(DASHB_EINVAL, "dashb_projection_static: empty or oversized problem");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return report_error(DASHB_ECUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (initialised_device() < 0 || dev != initialised_device() || dev >= MAX_DEVICES)
        return report_error(DASHB_ENODEV, "dashb_projection_static: call dashb_init() on the current device first");
    const ModInfo mi = modinfo_for((uint32_t)q_in), mo = modinfo_for((uint32_t)q_out);
    const uint64_t Bp = (batch + 15) & ~15ull;
    const uint64_t in_bytes = n * ((mi.nc + 3) / 4) * 4 * Bp;
    const uint64_t out_bytes = n * ((mo.nc + 3) / 4) * 4 * Bp;
    if (in_bytes + out_bytes > DASHB_STATIC_SCRATCH_BYTES)
        return report_error(DASHB_EINVAL,
                            "dashb_projection_static: %llu gates x %llu inferences need %llu bytes of digit planes, "
                            "the static scratch holds %u; split the call",
                            (unsigned long long)n, (unsigned long long)batch,
                            (unsigned long long)(in_bytes + out_bytes), (unsigned)DASHB_STATIC_SCRATCH_BYTES);
    unsigned char* scratch = nullptr;
    e = static_scratch_address(dev, &scratch);
    if (e != cudaSuccess) return report_error(DASHB_ECUDA, "cudaGetSymbolAddress(static scratch): %s", cudaGetErrorString(e));
    cudaStream_t s = stream ? (cudaStream_t)stream : library_stream();
    unsigned char* plane_in = scratch;
    unsigned char* plane_out = scratch + in_bytes; /* in_bytes is a multiple of 64 */
    /* the array still holds the previous call's planes: start from zeros */
    e = cudaMemsetAsync(scratch, 0, in_bytes + out_bytes, s);
    if (e == cudaSuccess) e = launch_ingest(d_in, plane_in, n, mi.nc, (uint32_t)batch, (uint32_t)Bp, s);
    if (e == cudaSuccess) {
        ProjParams p;
        memset(&p, 0, sizeof(p));
        p.k = 1;
        p.n = n;
        p.batch = (uint32_t)batch;
        p.Bp = (uint32_t)Bp;
        p.in[0] = plane_in;
        p.out[0] = plane_out;
        p.min[0] = mi;
        p.mout[0] = mo;
        p.ciphers[0] = reinterpret_cast<const gc::u128*>(d_ciphers);
        e = launch_projection(p, s);
    }
    if (e == cudaSuccess) e = launch_egress(plane_out, d_out, n, mo.nc, (uint32_t)batch, (uint32_t)Bp, s);
    if (e != cudaSuccess) return report_error(DASHB_ECUDA, "dashb_projection_static: %s", cudaGetErrorString(e));
    return DASHB_OK;
}

namespace dashb {

/* ------------------------------------------------- table replication */
/* out[r * rep + i] = in[r]: the tables of `rep` garblings interleaved row by
 * row (here `rep` copies of one garbling: same results, the memory traffic
 * of one garbling per inference; see SignParams::rep) */
__global__ void k_replicate_rows(const uint4* __restrict__ in, uint4* __restrict__ out, uint64_t rows, uint32_t rep) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < rows * rep; i += (uint64_t)gridDim.x * blockDim.x)
        out[i] = in[i / rep];
}
cudaError_t launch_replicate_rows(const void* in, void* out, uint64_t rows, uint32_t rep, cudaStream_t s) {
    if (rows == 0) return cudaSuccess;
    k_replicate_rows<<<148 * 8, 256, 0, s>>>((const uint4*)in, (uint4*)out, rows, rep);
    return cudaGetLastError();
}

/* ------------------------------------ output-channel sharding exchange */
/*
 * The all-gather of output-channel sharding as plain stores into peer
 * memory: after a layer group every rank owns a block of elements of the
 * activation tensor; k_shard_push reads that block once and stores it into
 * the gather buffer of EVERY rank (NVLink / NVSwitch posted writes for the
 * peers, a local store for itself).  k_shard_signal then publishes the lane's
 * epoch in every rank's flag array and k_shard_wait holds the lane's stream
 * until all ranks have published it -- the next layer reads a complete
 * tensor.  No NCCL, no host synchronisation; with two sub-batch lanes one
 * lane's exchange overlaps the other lane's layer group.  Gather buffers are
 * double-buffered per lane: a rank can only be pushing group g+2 into the
 * buffer group g used once every rank has finished group g+1, i.e. stopped
 * reading it.  (Plain loads and stores only: no TMA, no multicast.)
 */
__global__ void __launch_bounds__(256) k_shard_push(const __grid_constant__ ShardPushParams P) {
    const int j = blockIdx.y;
    /* the slice: nseg runs of seg elements (one run: the contiguous block [e0, e1)) */
    const uint64_t seg = P.seg ? P.seg : P.e1 - P.e0, nseg = P.seg ? P.nseg : 1;
    const uint64_t vrun = seg * P.row[j] / 16; /* rows are multiples of 64 bytes */
    const uint64_t nvec = vrun * nseg;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t run = i / vrun, w = i - run * vrun;
        const uint64_t at = (P.e0 + run * P.stride) * P.row[j] + 16 * w; /* byte offset inside the plane */
        const uint4 v = *reinterpret_cast<const uint4*>(P.src[j] + at);
        for (int r = 0; r < P.world; ++r) *reinterpret_cast<uint4*>(P.peer[r] + P.dst_off[j] + at) = v;
    }
    __threadfence_system();
}

struct ShardFlagPtrs {
    uint32_t* p[MAXR];
};
__global__ void k_shard_signal(const ShardFlagPtrs F, int world, int rank, uint32_t epoch) {
    if ((int)threadIdx.x < world) {
        __threadfence_system();
        *reinterpret_cast<volatile uint32_t*>(F.p[threadIdx.x] + rank) = epoch;
        __threadfence_system();
    }
}
/* flags[r] >= epoch for every rank r; gives up after ~4 s of spinning (a
 * peer that died must not hang this GPU) and records that in flags[MAXR] */
__global__ void k_shard_wait(uint32_t* flags, int world, uint32_t epoch) {
    if ((int)threadIdx.x < world) {
        const long long t0 = clock64();
        while ((int32_t)(*reinterpret_cast<volatile uint32_t*>(flags + threadIdx.x) - epoch) < 0) {
            if (clock64() - t0 > 8000000000ll) {
                *reinterpret_cast<volatile uint32_t*>(flags + MAXR) = epoch;
                break;
            }
            __nanosleep(200);
        }
        __threadfence_system();
    }
}

cudaError_t launch_shard_push(const ShardPushParams& p, cudaStream_t s) {
    const uint64_t count = p.seg ? p.seg * p.nseg : (p.e1 > p.e0 ? p.e1 - p.e0 : 0);
    if (count == 0) return cudaSuccess;
    uint64_t nvec = 0;
    for (int j = 0; j < p.k; ++j) nvec = std::max<uint64_t>(nvec, count * p.row[j] / 16);
    /* a modest grid: the exchange shares the GPU with the other lane's layer group */
    static const int per_sm = getenv("DASHB_PUSH_CTAS_PER_SM") ? atoi(getenv("DASHB_PUSH_CTAS_PER_SM")) : 2;
    dim3 grid((unsigned)std::min<uint64_t>((nvec + 255) / 256, (uint64_t)148 * std::max(1, per_sm)), p.k);
    k_shard_push<<<grid, 256, 0, s>>>(p);
    return cudaGetLastError();
}
cudaError_t launch_shard_signal(uint32_t* const* peer_flags, int world, int rank, uint32_t epoch, cudaStream_t s) {
    ShardFlagPtrs f;
    memset(&f, 0, sizeof(f));
    for (int r = 0; r < world; ++r) f.p[r] = peer_flags[r];
    k_shard_signal<<<1, 32, 0, s>>>(f, world, rank, epoch);
    return cudaGetLastError();
}
cudaError_t launch_shard_wait(uint32_t* flags, int world, uint32_t epoch, cudaStream_t s) {
    k_shard_wait<<<1, 32, 0, s>>>(flags, world, epoch);
    return cudaGetLastError();
}

/* --------------------------------------------------------- max pooling */
/*
 * GarbledMaxPool2d::cpu_evaluate (layer/garbled_maxpool2d.h:143-224): every
 * output (i, j, channel) starts from the window's first input and takes, for
 * every kernel position (l, m) in turn -- (0, 0) included -- the MaxGadget
 * step  cur = cur + relu(in(i s_w + l, j s_h + m) - cur)
 * (gadgets/max_gadget.h:116-160), each step with a ReLU gadget of its own
 * (gadgets/relu_gadget.h).  The ReLU is k_sign<relu> with the tables of gadget
 * instance  output * (k_w k_h) + (l k_h + m);  this kernel is the label
 * arithmetic between two of them, one 32-bit word (four digits of one
 * inference) per thread.  The layer addresses labels with the last index
 * fastest on both sides: input (x, y, z) at (x H + y) C + z
 * (LabelSlice::get_first_label_id_from_slice_dims, label_tensor.h:1574-1579),
 * output (i, j, channel) at (i oh + j) C + channel (set_label, :1290-1301).
 */
__device__ __forceinline__ uint32_t bytes_addsub(uint32_t a, uint32_t b, uint32_t q, bool sub) {
    uint32_t r = 0;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const uint32_t x = (a >> (8 * t)) & 0xff, y = (b >> (8 * t)) & 0xff;
        uint32_t v = sub ? x + q - y : x + y;
        v = v >= q ? v - q : v;
        r |= v << (8 * t);
    }
    return r;
}

__global__ void __launch_bounds__(256) k_pool_step(const __grid_constant__ PoolParams P) {
    const int j = blockIdx.y;
    const uint32_t q = P.crt[j].q;
    const uint64_t per = (uint64_t)((P.crt[j].nc + 3) >> 2) * P.Bp; /* words of one element */
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= P.n_out * per) return;
    const uint64_t o = w / per, rest = w - o * per;
    /* this layer counts the LAST index fastest on both sides (slices: label_tensor.h:1574-1579, set_label: :1290-1301) */
    const uint32_t z = (uint32_t)(o % P.ch), jj = (uint32_t)((o / P.ch) % P.oh), i = (uint32_t)(o / ((uint64_t)P.ch * P.oh));
    const uint32_t* in = reinterpret_cast<const uint32_t*>(P.in[j]);
    uint32_t* cur = reinterpret_cast<uint32_t*>(P.cur[j]);
    uint32_t c;
    if (P.mode == 0) {
        const uint64_t first = ((uint64_t)(i * P.s_w) * P.in_h + jj * P.s_h) * P.ch + z;
        c = in[first * per + rest];
    } else {
        c = bytes_addsub(cur[w], reinterpret_cast<const uint32_t*>(P.relu[j])[w], q, false);
    }
    cur[w] = c;
    if (P.mode == 2) return;
    const uint64_t pos = ((uint64_t)(i * P.s_w + P.l) * P.in_h + (jj * P.s_h + P.m)) * P.ch + z;
    reinterpret_cast<uint32_t*>(P.diff[j])[w] = bytes_addsub(in[pos * per + rest], c, q, true);
}

cudaError_t launch_pool_step(const PoolParams& p, cudaStream_t s) {
    if (p.n_out == 0) return cudaSuccess;
    uint64_t words = 0;
    for (int j = 0; j < p.k; ++j) words = std::max<uint64_t>(words, p.n_out * ((p.crt[j].nc + 3) >> 2) * p.Bp);
    dim3 grid((unsigned)((words + 255) / 256), p.k);
    k_pool_step<<<grid, 256, 0, s>>>(p);
    return cudaGetLastError();
}

/* ------------------------------------------------- sign gadget / ReLU */
/* Per-thread digit accumulators live in shared memory, digit-major
 * (acc[d * blockDim + tid]) so that a warp's accesses are conflict-free. */
#ifndef DASHB_SIGN_THREADS
#define DASHB_SIGN_THREADS 128
#endif
constexpr int SIGN_THREADS = DASHB_SIGN_THREADS;
#ifndef DASHB_SIGN_MINB
#define DASHB_SIGN_MINB 8 /* CTAs per SM the register allocation aims at */
#endif

/* Digit accumulators of one thread: 32-bit words holding two digits' sums
 * (digit 2w in the low half, 2w+1 in the high half), word-major in shared
 * memory (acc[w * SIGN_THREADS]).  A sum of k+1 digits stays below 2^16
 * (checked when the circuit is created), so the halves never meet. */

/* Digit-wise accumulation of a compressed label by long division, least
 * significant digit first: the fallback for a modulus without a
 * multiplicative plan.  One copy per kernel. */
__device__ __noinline__ void accumulate(u128 x, const ModInfo m, uint32_t* acc) {
    gc::for_each_digit(x, m, [&](uint32_t d, uint32_t v) { acc[(d >> 1) * SIGN_THREADS] += v << (16 * (d & 1)); });
}

/* The same accumulation with the digits taken by multiplication, most
 * significant chunk first (gc::for_each_chunk_ladder; labels of a correct
 * evaluation are always below q^nc).  STAGE-LOCAL PAIR LAYOUT: the chunk of
 * stage s (len_s digits) owns ceil(len_s / 2) accumulator words of its own,
 * digit pair (2w, 2w+1) of the chunk in word w -- chunks never share a word,
 * whatever the parity of their lengths, so the inner loop is two reciprocal
 * multiplies, two multiply-adds that also pack the pair (d0 + 2^16 d1 =
 * r + q1 (2^16 - q) - q2 (q 2^16)) and one shared-memory read-modify-write
 * per pair, nothing else.  compress_acc_ladder reads the same layout. */
#ifndef DASHB_FRAC
#define DASHB_FRAC 1
#endif
__device__ __forceinline__ void accumulate_ladder(u128 x, const ModInfo& m, const gc::FracInfo& fi, uint32_t* acc) {
    /* c2 = -(q << 16): the subtraction rides on the multiply-add */
    const uint32_t c1 = 65536u - m.q, c2 = 0u - (m.q << 16), mq = m.mq;
    uint32_t* a = acc;
    gc::for_each_chunk_ladder(x, fi, [&](uint32_t len, uint32_t r) {
#pragma unroll 1
        for (; len >= 4; len -= 4) { /* two words per trip */
            const uint32_t q1 = gc::mulhi32(r, mq);
            const uint32_t q2 = gc::mulhi32(q1, mq);
            const uint32_t q3 = gc::mulhi32(q2, mq);
            const uint32_t q4 = gc::mulhi32(q3, mq);
            a[0] += r + q1 * c1 + q2 * c2;
            a[SIGN_THREADS] += q2 + q3 * c1 + q4 * c2;
            a += 2 * SIGN_THREADS;
            r = q4;
        }
        if (len & 2) {
            const uint32_t q1 = gc::mulhi32(r, mq);
            const uint32_t q2 = gc::mulhi32(q1, mq);
            a[0] += r + q1 * c1 + q2 * c2;
            a += SIGN_THREADS;
            r = q2;
        }
        if (len & 1) { /* the chunk's last digit is what is left */
            a[0] += r;
            a += SIGN_THREADS;
        }
    });
}

/* The shape nearly every cast modulus has (FracInfo::t4: five or six stages
 * of four digits, widths 5,[5,]4,3,2,1) as straight-line code: no stage loop,
 * no length dispatch, accumulator words at fixed offsets.  The last of six
 * stages may hold fewer than four digits; it is run as four all the same (its
 * missing digits come out zero), which can touch one word past the layout:
 * launch_sign sizes the accumulators with that word of slack. */
__device__ __forceinline__ void accumulate_t4(u128 x, const ModInfo& m, const gc::FracInfo& fi, uint32_t* acc) {
    uint32_t n[5];
    gc::frac_init(x, fi, n);
    const uint32_t c1 = 65536u - m.q, c2 = 0u - (m.q << 16), mq = m.mq;
    const uint32_t multm = fi.mult[1];
    auto quad = [&](uint32_t r, uint32_t* a) {
        const uint32_t q1 = gc::mulhi32(r, mq);
        const uint32_t q2 = gc::mulhi32(q1, mq);
        const uint32_t q3 = gc::mulhi32(q2, mq);
        const uint32_t q4 = gc::mulhi32(q3, mq);
        a[0] += r + q1 * c1 + q2 * c2;
        a[SIGN_THREADS] += q2 + q3 * c1 + q4 * c2;
    };
    uint32_t* a = acc;
    quad(gc::frac_stage(n, fi.mult[0], 5), a);
    if (fi.nst == 6) {
        a += 2 * SIGN_THREADS;
        quad(gc::frac_stage(n, multm, 5), a);
    }
    quad(gc::frac_stage(n, multm, 4), a + 2 * SIGN_THREADS);
    quad(gc::frac_stage(n, multm, 3), a + 4 * SIGN_THREADS);
    quad(gc::frac_stage(n, multm, 2), a + 6 * SIGN_THREADS);
    quad(gc::frac_stage(n, fi.mult[fi.nst - 1], 1), a + 8 * SIGN_THREADS);
}

/* compress(acc mod q) and the colour of the sum; absolute digit layout
 * (accumulate) */
/* T > 1: the T threads of an element each hold a partial sum in their own
 * accumulator column (adjacent words, acc[.. + 0 .. T-1]); the reader adds
 * them up -- packed halves still cannot meet: the k + 1 summands are the same */
template <int T>
__device__ __forceinline__ uint32_t acc_word(const uint32_t* a) {
    uint32_t v = a[0];
#pragma unroll
    for (int u = 1; u < T; ++u) v += a[u];
    return v;
}

template <int T>
__device__ __forceinline__ u128 compress_acc(const uint32_t* acc, const ModInfo& m, uint32_t& colour) {
    gc::Compressor c(m);
    uint32_t dg = 0;
    for (int d = (int)m.nc - 1; d >= 0; --d) {
        const uint32_t w = acc_word<T>(acc + (d >> 1) * SIGN_THREADS);
        dg = gc::mod_q((d & 1) ? (w >> 16) : (w & 0xffffu), m);
        c.push(dg);
    }
    colour = dg;
    return c.value();
}

/* the same for the stage-local pair layout (accumulate_ladder): stage after
 * stage from the most significant one, a stage's pairs from its top word
 * down; a stage's digits are folded into one 32-bit value (below q^len_s =
 * the stage's multiplier, 2^sh removed for stage 0) that enters the 128-bit
 * value with one multiply-add.  An odd stage's top word has an empty high
 * half: a leading zero. */
template <int T>
__device__ __forceinline__ u128 compress_acc_ladder(const uint32_t* acc, const ModInfo& m, const gc::FracInfo& fi,
                                                    uint32_t& colour) {
    u128 x = gc::make_u128(0, 0);
    const uint32_t q = m.q, mq = m.mq;
    uint32_t lo = 0;
#pragma unroll 1
    for (uint32_t s = 0; s < fi.nst; ++s) {
        const uint32_t nw = ((fi.desc[s] & 0xffu) + 1) >> 1;
        uint32_t chunk = 0;
#pragma unroll 1
        for (int w = (int)nw - 1; w >= 0; --w) {
            const uint32_t v = acc_word<T>(acc + w * SIGN_THREADS);
            const uint32_t hi = v >> 16;
            lo = v & 0xffffu;
            chunk = (chunk * q + (hi - gc::mulhi32(hi, mq) * q)) * q;
            lo -= gc::mulhi32(lo, mq) * q;
            chunk += lo;
        }
        /* q^len_s: stage 0's multiplier carries the 2^sh alignment factor, but x is still 0 there */
        x = gc::muladd_small(x, fi.mult[s], chunk);
        acc += nw * SIGN_THREADS;
    }
    colour = lo;
    return x;
}

/* ReLU step 4 for one CRT modulus with a word-aligned plan: out = E + t x - G
 * digit-wise mod p (gates/mixed_mod_half_gate.h:100-102), the digits of the
 * two payloads taken by multiplication in lockstep.  Every stage of the plan
 * covers whole 32-bit words of the digit plane (make_fracinfo(aligned4));
 * the first one is rounded up to a whole word: the payloads have no digits
 * above nc and the input label's padding bytes are masked, so the padding of
 * the output word comes out zero.  Per word: 8 reciprocal multiplies and 8
 * multiply-adds give the packed digit pairs of E and G, two PRMT spread the
 * four x digits, then  pair = E + t x + (q,q) - G  is reduced half by half. */
__device__ __forceinline__ void half_gate_ladder(u128 E, u128 G, const ModInfo& m, const gc::FracInfo& fi, uint32_t tt,
                                                 const uint8_t* xin, uint8_t* xout, uint32_t Bp) {
    uint32_t ne[5], ng[5];
    gc::frac_init(E, fi, ne);
    gc::frac_init(G, fi, ng);
    const uint32_t q = m.q, mq = m.mq, c1 = 65536u - q, c2 = 0u - (q << 16), qq = q | (q << 16), nq = 0u - q;
    const uint32_t topmask = 0xffffffffu >> (8 * ((4u - (m.nc & 3u)) & 3u));
    uint32_t wstride32 = 4u * Bp;
    asm volatile("" : "+r"(wstride32)); /* keep the stride in a register (else re-read from the parameter bank per word) */
    const uint64_t wstride = wstride32;
    uint32_t d = (m.nc + 3u) & ~3u; /* digits not yet delivered, rounded up to whole words */
    uint32_t s = 0;
    const uint32_t cnt = fi.ladder;
    bool first = true; /* only a stage of the widest loop can be the first one */
    auto stage = [&](uint32_t len, uint32_t re, uint32_t rg, bool first) {
        const uint32_t nw = (len + 3u) >> 2;
        d -= 4 * nw;
        const uint8_t* xp = xin + (uint64_t)(d >> 2) * wstride;
        uint8_t* op = xout + (uint64_t)(d >> 2) * wstride;
        uint32_t xn = *reinterpret_cast<const uint32_t*>(xp);
#pragma unroll 1
        for (uint32_t w = 0; w < nw; ++w) {
            uint32_t xw = xn;
            if (w + 1 < nw) xn = *reinterpret_cast<const uint32_t*>(xp + wstride);
            xp += wstride;
            if (first && w + 1 == nw) xw &= topmask;
            const uint32_t e1 = gc::mulhi32(re, mq), e2 = gc::mulhi32(e1, mq), e3 = gc::mulhi32(e2, mq), e4 = gc::mulhi32(e3, mq);
            const uint32_t g1 = gc::mulhi32(rg, mq), g2 = gc::mulhi32(g1, mq), g3 = gc::mulhi32(g2, mq), g4 = gc::mulhi32(g3, mq);
            const uint32_t pe0 = re + e1 * c1 + e2 * c2, pe1 = e2 + e3 * c1 + e4 * c2;
            const uint32_t pg0 = rg + g1 * c1 + g2 * c2, pg1 = g2 + g3 * c1 + g4 * c2;
            re = e4;
            rg = g4;
            const uint32_t v0 = tt * __byte_perm(xw, 0, 0x4140) + pe0 + qq - pg0;
            const uint32_t v1 = tt * __byte_perm(xw, 0, 0x4342) + pe1 + qq - pg1;
            uint32_t a0 = v0 & 0xffffu, a1 = v0 >> 16, a2 = v1 & 0xffffu, a3 = v1 >> 16;
            a0 += gc::mulhi32(a0, mq) * nq;
            a1 += gc::mulhi32(a1, mq) * nq;
            a2 += gc::mulhi32(a2, mq) * nq;
            a3 += gc::mulhi32(a3, mq) * nq;
            *reinterpret_cast<uint32_t*>(op) = a0 | (a1 << 8) | (a2 << 16) | (a3 << 24);
            op += wstride;
        }
    };
#define DASHB_HG_STEP(W)                                                        \
    _Pragma("unroll 1") for (uint32_t c = (cnt >> (4 * ((W)-1))) & 15u; c; --c, ++s) { \
        const uint32_t mult = fi.mult[s];                                         \
        const uint32_t re = gc::frac_stage(ne, mult, (W));                        \
        const uint32_t rg = gc::frac_stage(ng, mult, (W));                        \
        stage(fi.desc[s] & 0xffu, re, rg, (W) == 5 && first);                     \
        first = false;                                                            \
    }
    DASHB_HG_STEP(5)
    DASHB_HG_STEP(4)
    DASHB_HG_STEP(3)
    DASHB_HG_STEP(2)
    DASHB_HG_STEP(1)
#undef DASHB_HG_STEP
}

/*
 * Fused approximate-sign gadget, optionally followed by the ReLU's mixed-
 * modulus half gates.  Replaces eval_approx_res_gadget + mrs_sum_most_sig +
 * eval_sign (gadgets/sign_gadget.h:41-213,737-795) and eval_sign_mult
 * (layer/garbled_relu.h:19-38,258-294) -- three kernels, three device
 * syncs, device-side new/delete -- and follows SignGadget::cpu_evaluate
 * (gadgets/sign_gadget.h:583-671) + MixedModHalfGate::cpu_evaluate
 * (gates/mixed_mod_half_gate.h:91-104) bit for bit.
 *
 * Hash reuse (results identical, fewer AES blocks than the reference):
 * the n_mrs approximate-residue projections of residue j and its garbler
 * half gate all key on H(x_j); the evaluator half gate and the mini gate of
 * every modulus key on H(s); the sign projections of every out modulus key
 * on H(top sum).  The per-digit sums of digits i >= 1 (mrs_sum, :615-622)
 * are never read by the reference and are not formed.
 */
/* T threads per (element, inference), T = 1 or 4.  With T = 4 the residues
 * are dealt round the four threads of a quad (thread t takes j = t, t+4, ...;
 * the carry counts as residue k): the input hashes, the cast-1 projections of
 * every mixed-radix digit and the half gates run four abreast, each thread
 * accumulating into its own column; thread 0 adds the four columns up while it
 * compresses, does the one cast-2 hash of the digit and hands the carry back
 * by shuffle.  The dependent chain of an element shrinks from k + (k+2)(n-1)
 * + 3 hashes to about a third, for ~1.4x the issue slots: launch_sign uses it
 * for launches too small to fill the GPU, where the chain's latency -- not
 * throughput -- is what the layer costs (the small ReLU layers of a CNN, a
 * rank's share of a channel-sharded layer, a small batch). */
template <bool RELU, int T>
__global__ void __launch_bounds__(SIGN_THREADS, DASHB_SIGN_MINB) k_sign(const __grid_constant__ SignParams P) {
    __shared__ AesSmem aes;
    extern __shared__ uint32_t s_acc[]; /* [accumulator words][SIGN_THREADS] */
    aes_smem_load(aes);
    const uint32_t t = T > 1 ? (threadIdx.x % T) : 0u;
    const uint32_t qmask = T > 1 ? (((1u << T) - 1u) << ((threadIdx.x & 31u) & ~(uint32_t)(T - 1))) : 0xffffffffu; /* the element's threads */
    const uint64_t idx = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / T;
    if (idx >= P.n * P.batch) return; /* whole quads leave together */
    const uint32_t b = (uint32_t)(idx % P.batch);
    const uint64_t e = slice_elem(P.e0, P.seg, P.stride, idx / P.batch);
    const int k = P.k;
    const int nm = P.n_mrs;

    /* Step 1a: key and hash every CRT residue label once */
    u128 h[(MAXK + T - 1) / T]; /* residue j lives at [j / T] of thread j % T */
    uint32_t col[(MAXK + T - 1) / T];
#pragma unroll 1
    for (int j = t; j < k; j += T) {
        const ModInfo Pm = pin(P.crt[j]);
        const u128 key = load_compress(P.in[j], Pm, e, b, P.Bp, col[j / T]);
        h[j / T] = hash(aes, key);
    }

    /* Step 2: mixed-radix addition from the least significant digit */
    /* hide the thread index's origin: under register pressure the compiler
     * otherwise re-reads %tid.x (S2R) inside the digit loops to rebuild the
     * accumulator address */
    uint32_t tid = threadIdx.x;
    asm volatile("" : "+r"(tid));
    uint32_t* acc = s_acc + tid;
    u128 carry = P.carry0;
    /* rows of this element's tables, as 32-bit row indices from the element's
     * first row (a table of one element never reaches 2^32 rows) */
    const uint64_t et = e * P.tab_mul + P.tab_add; /* the gadget instance whose tables this element uses */
    /* table replicas (SignParams::rep): row r of this inference sits at r * R + rb */
    const uint32_t R = P.rep;
    const uint32_t rb = R > 1 ? (P.b0 + b) % R : 0u;
    const u128* approx_e = P.approx + (et * (uint64_t)nm * P.S) * R + rb;
    const u128* cast1_e = P.cast1 + (et * (uint64_t)(k + 1) * P.Mp) * R + rb;
    const u128* cast2_e = P.cast2 + (et * (uint64_t)(k + 1) * P.Mp) * R + rb;
    uint32_t cnt1 = 0, cnt2 = 0;
    uint32_t ctop = 0;
    u128 top = gc::make_u128(0, 0);
    /* One loop over the mixed-radix digits, least significant first; digit 0
     * (the most significant) skips the cast: its k + 1 summands are the
     * approximate digits themselves and the carry. */
#pragma unroll 1
    for (int i = nm - 1; i >= 0; --i) {
        /* by value: keeps the fields in registers instead of re-reading the
         * dynamically indexed parameter bank at every use */
        const ModInfo M = pin(P.mrs[i]);
        const ModInfo C = pin(P.cast[i]); /* accumulation modulus: (k+1) m_i, or m_0 for digit 0 */
        const gc::ColourConsts cc = gc::colour_consts(M);
        /* this digit's plan, addressed through an offset the compiler cannot
         * rebuild from i: otherwise every stage of every label recomputes
         * &P.castf[i] (six uniform-datapath instructions) before its loads */
        uint32_t plan_off = (uint32_t)(offsetof(SignParams, castf) + (size_t)i * sizeof(gc::FracInfo));
        asm volatile("" : "+r"(plan_off));
        const gc::FracInfo& fi = *reinterpret_cast<const gc::FracInfo*>(reinterpret_cast<const char*>(&P) + plan_off);
        const bool ladder = DASHB_FRAC && fi.nst != 0;
        const bool t4 = ladder && fi.t4 != 0;
        const uint32_t nwords = ladder ? fi.words : (C.nc + 1) / 2;
        for (uint32_t w = 0; w < nwords; ++w) acc[w * SIGN_THREADS] = 0;
        /* Ciphertext rows are requested ahead of the hashes that separate
         * them from their use: residue j+1's approximate-digit row before
         * label j is hashed, the cast-1 row (its index is the label's colour,
         * known before the hash) before the hash as well. */
        auto approx_row = [&](int j) { return ldg128(approx_e + (uint64_t)((uint32_t)nm * P.prefix[j] + (uint32_t)i * P.crt[j].q + col[j / T]) * R); };
        u128 row = gc::make_u128(0, 0);
        if ((int)t < k) row = approx_row(t);
#pragma unroll 1
        for (int j = t; j <= k; j += T) {
            /* Step 1b: approx. MRS digit i of residue j (j == k: the carry) */
            u128 lab = carry;
            if (j < k) lab = gc::sub128(row, h[j / T]);
            if (j + T < k) row = approx_row(j + T);
            if (i > 0) {
                /* Step 2.2: cast to modulus (k+1)*m_i */
                const uint32_t c0 = gc::colour_of(lab, M, cc);
                const u128 crow = ldg128(cast1_e + (uint64_t)(cnt1 + (uint32_t)j * M.q + c0) * R);
                lab = gc::sub128(crow, hash(aes, lab));
            }
            if (t4)
                accumulate_t4(lab, C, fi, acc);
            else if (ladder)
                accumulate_ladder(lab, C, fi, acc);
            else
                accumulate(lab, C, acc);
        }
        /* Step 2.3 / 2.4: sum of the casts -> carry into digit i-1 (thread 0 of the element) */
        if (T > 1) __syncwarp(qmask); /* every column of the element is complete */
        if (t == 0) {
            uint32_t c2;
            const u128 sum = ladder ? compress_acc_ladder<T>(acc, C, fi, c2) : compress_acc<T>(acc, C, c2);
            if (i > 0) {
                carry = gc::sub128(ldg128(cast2_e + (uint64_t)(cnt2 + c2) * R), hash(aes, sum));
            } else {
                top = sum;
                ctop = c2;
            }
        }
        if (i > 0) {
            cnt1 += (uint32_t)(k + 1) * M.q;
            cnt2 += C.q;
        }
        if (T > 1) { /* the carry goes to the thread that owns residue k; also: columns may be cleared again */
            carry.lo = __shfl_sync(qmask, carry.lo, 0, T);
            carry.hi = __shfl_sync(qmask, carry.hi, 0, T);
        }
    }
    u128 htop = gc::make_u128(0, 0);
    if (t == 0) htop = hash(aes, top);
    if (T > 1) {
        htop.lo = __shfl_sync(qmask, htop.lo, 0, T);
        htop.hi = __shfl_sync(qmask, htop.hi, 0, T);
        ctop = __shfl_sync(qmask, ctop, 0, T);
    }
    const uint32_t m0 = P.mrs[0].q;

    /* Step 3: sign projection(s) */
    const u128* sign_e = P.sign + (et * (uint64_t)P.n_out * m0) * R + rb;
    if (!RELU) {
#pragma unroll 1
        for (int j = t; j < P.n_out; j += T) {
            const ModInfo mo = pin(P.outm[j]);
            store_digits(P.out[j], mo, e, b, P.Bp, gc::sub128(ldg128(sign_e + ((uint64_t)j * m0 + ctop) * R), htop));
        }
        return;
    }
    /* Step 4 (ReLU): out_j = E_j + (y+r)*x_j - G_j  (mod p_j) */
    const u128 s = gc::sub128(ldg128(sign_e + (uint64_t)ctop * R), htop); /* sign label, modulus 2 */
    u128 hs = gc::make_u128(0, 0);
    if (t == 0) hs = hash(aes, s);
    if (T > 1) {
        hs.lo = __shfl_sync(qmask, hs.lo, 0, T);
        hs.hi = __shfl_sync(qmask, hs.hi, 0, T);
    }
    const uint32_t cs = (uint32_t)s.lo & 1u;
    const u128* garbler_e = P.garbler + (et * (uint64_t)P.S) * R + rb;
    const u128* ev_e = P.evaluator + (et * (uint64_t)k * 3) * R + rb;
#pragma unroll 1
    for (int j = t; j < k; j += T) {
        const ModInfo Pm = pin(P.crt[j]);
        const u128 G = gc::sub128(ldg128(garbler_e + (uint64_t)(P.prefix[j] + col[j / T]) * R), h[j / T]);
        const u128* ev = ev_e + (uint64_t)(3 * j) * R;
        const u128 E = gc::sub128(ldg128(ev + (uint64_t)cs * R), hs);
        /* mini gate: int16 row `cs` of the packed slot minus the low 16 hash
         * bits (gates/projection_gate_mini.h:45-56) */
        const u128 mini = ldg128(ev + (uint64_t)2 * R);
        const int16_t row16 = (int16_t)((mini.lo >> (16 * cs)) & 0xffff);
        const int16_t t16 = (int16_t)(row16 - (int16_t)(hs.lo & 0xffff));
        int tt = (int)t16 % (int)Pm.q;
        if (tt < 0) tt += (int)Pm.q;
        const uint64_t off = plane_off(Pm, e, b, P.Bp);
        const uint8_t* xin = P.in[j] + off;
        uint8_t* xout = P.out[j] + off;
        if (DASHB_FRAC && P.crtf[j].nst) {
            half_gate_ladder(E, G, Pm, P.crtf[j], (uint32_t)tt, xin, xout, P.Bp);
            continue;
        }
        uint32_t xw = 0, ow = 0;
        const uint32_t last = Pm.nc - 1;
        /* the input label's words again (they were read for the key in step
         * 1a), two ahead of their use */
        const uint32_t nw = nc4_of(Pm);
        auto xword = [&](uint32_t d4) { return d4 < nw ? *reinterpret_cast<const uint32_t*>(xin + (uint64_t)d4 * 4 * P.Bp) : 0u; };
        uint32_t xn0 = xword(0), xn1 = xword(1);
        gc::for_each_digit2(E, G, Pm, [&](uint32_t d, uint32_t de, uint32_t dg) {
            if ((d & 3) == 0) {
                xw = xn0;
                xn0 = xn1;
                xn1 = xword((d >> 2) + 2);
            }
            const uint32_t x = (xw >> (8 * (d & 3))) & 0xff;
            ow |= gc::mod_q(de + (uint32_t)tt * x + Pm.q - dg, Pm) << (8 * (d & 3));
            if ((d & 3) == 3 || d == last) {
                *reinterpret_cast<uint32_t*>(xout + (uint64_t)(d >> 2) * 4 * P.Bp) = ow;
                ow = 0;
            }
        });
    }
}

/* ------------------------------------------ sign gadget, latency mode */
/*
 * Same gadget, one WARP per (element, inference): for small launches (a
 * single inference, the small layers of a small batch) k_sign leaves most of
 * the GPU idle while every thread walks its ~88 hashes one after the other.
 * Here the k input hashes, the k*(n-1) cast-1 projections and the k top-digit
 * projections are spread over the lanes (they do not depend on the carry);
 * only the carry chain itself -- two hashes per mixed-radix digit -- stays
 * sequential on lane 0, and the half gates again run one modulus per lane.
 * Critical path: ~2n+4 hashes instead of k*n + 2n.  It issues ~7x the
 * instructions per element, so launch_sign only picks it when the launch
 * could not fill the machine anyway.  Results identical to k_sign.
 */
constexpr int WIDE_WARPS = 4;

/* acc[d] += digit_d(x), shared-memory atomics (several lanes add into the same digits) */
__device__ __noinline__ void accumulate_atomic(u128 x, const ModInfo m, uint32_t* acc) {
    gc::for_each_digit(x, m, [&](uint32_t d, uint32_t v) { atomicAdd(acc + d, v); });
}

__device__ __forceinline__ u128 compress_acc32(const uint32_t* acc, const ModInfo& m, uint32_t& colour) {
    gc::Compressor c(m);
    uint32_t dg = 0;
    for (int d = (int)m.nc - 1; d >= 0; --d) {
        dg = gc::mod_q(acc[d], m);
        c.push(dg);
    }
    colour = dg;
    return c.value();
}

__device__ __forceinline__ u128 shfl128(u128 v, int src) {
    u128 r;
    r.lo = __shfl_sync(0xffffffffu, v.lo, src);
    r.hi = __shfl_sync(0xffffffffu, v.hi, src);
    return r;
}

template <bool RELU>
__global__ void __launch_bounds__(WIDE_WARPS * 32) k_sign_wide(const __grid_constant__ SignParams P, uint32_t acc_digits) {
    __shared__ AesSmem aes;
    __shared__ u128 s_h[WIDE_WARPS][MAXK];
    __shared__ uint32_t s_col[WIDE_WARPS][MAXK];
    extern __shared__ uint32_t s_wacc[]; /* [warp][n_mrs][acc_digits] */
    aes_smem_load(aes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint64_t idx = (uint64_t)blockIdx.x * WIDE_WARPS + warp;
    if (idx >= P.n * P.batch) return; /* whole warps leave together */
    const uint32_t b = (uint32_t)(idx % P.batch);
    const uint64_t e = slice_elem(P.e0, P.seg, P.stride, idx / P.batch);
    const int k = P.k, nm = P.n_mrs;
    u128* h = s_h[warp];
    uint32_t* col = s_col[warp];
    uint32_t* acc = s_wacc + (size_t)warp * nm * acc_digits;

    for (uint32_t t = lane; t < (uint32_t)nm * acc_digits; t += 32) acc[t] = 0;
    /* step 1a: hash every CRT residue label, one per lane */
    if (lane < k) {
        const ModInfo Pm = P.crt[lane];
        uint32_t c;
        const u128 key = load_compress(P.in[lane], Pm, e, b, P.Bp, c);
        h[lane] = hash(aes, key);
        col[lane] = c;
    }
    __syncwarp();
    const uint64_t et = e * P.tab_mul + P.tab_add; /* the gadget instance whose tables this element uses */
    const u128* approx_e = P.approx + et * (uint64_t)nm * P.S;
    const uint64_t cast_e = et * (uint64_t)(k + 1) * P.Mp;
    /* steps 1b + 2.2 for every (digit i >= 1, residue j): independent of the carry */
    for (int t = lane; t < k * (nm - 1); t += 32) {
        const int j = t % k, i = 1 + t / k;
        const ModInfo M = P.mrs[i];
        const ModInfo C = P.cast[i];
        uint64_t cnt1 = cast_e;
        for (int i2 = nm - 1; i2 > i; --i2) cnt1 += (uint64_t)(k + 1) * P.mrs[i2].q;
        const u128 lab = gc::sub128(ldg128(approx_e + (uint64_t)nm * P.prefix[j] + (uint64_t)i * P.crt[j].q + col[j]), h[j]);
        const u128 hl = hash(aes, lab);
        const uint32_t c0 = gc::colour_of(lab, M);
        accumulate_atomic(gc::sub128(ldg128(P.cast1 + cnt1 + (uint64_t)j * M.q + c0), hl), C, acc + (size_t)i * acc_digits);
    }
    /* most significant digit: the k approximate digits themselves */
    if (lane < k) {
        const u128 lab = gc::sub128(ldg128(approx_e + (uint64_t)nm * P.prefix[lane] + col[lane]), h[lane]);
        accumulate_atomic(lab, P.mrs[0], acc);
    }
    __syncwarp();
    /* the carry chain, least significant digit first (lane 0) */
    u128 htop = gc::make_u128(0, 0);
    uint32_t ctop = 0;
    if (lane == 0) {
        u128 carry = P.carry0;
        uint64_t cnt1 = cast_e, cnt2 = cast_e;
        for (int i = nm - 1; i >= 1; --i) {
            const ModInfo M = P.mrs[i];
            const ModInfo C = P.cast[i];
            uint32_t* a = acc + (size_t)i * acc_digits;
            const u128 hl = hash(aes, carry);
            const uint32_t c0 = gc::colour_of(carry, M);
            accumulate_atomic(gc::sub128(ldg128(P.cast1 + cnt1 + (uint64_t)k * M.q + c0), hl), C, a);
            cnt1 += (uint64_t)(k + 1) * M.q;
            uint32_t c2;
            const u128 sum2 = compress_acc32(a, C, c2);
            carry = gc::sub128(ldg128(P.cast2 + cnt2 + c2), hash(aes, sum2));
            cnt2 += C.q;
        }
        accumulate_atomic(carry, P.mrs[0], acc);
        const u128 top = compress_acc32(acc, P.mrs[0], ctop);
        htop = hash(aes, top);
    }
    htop = shfl128(htop, 0);
    ctop = __shfl_sync(0xffffffffu, ctop, 0);
    const uint32_t m0 = P.mrs[0].q;
    const u128* sign_e = P.sign + et * (uint64_t)P.n_out * m0;
    if (!RELU) {
        if (lane < P.n_out) store_digits(P.out[lane], P.outm[lane], e, b, P.Bp, gc::sub128(ldg128(sign_e + (uint64_t)lane * m0 + ctop), htop));
        return;
    }
    /* step 4 (ReLU): hash of the sign label on lane 0, then one modulus per lane */
    const u128 s = gc::sub128(ldg128(sign_e + ctop), htop);
    u128 hs = gc::make_u128(0, 0);
    if (lane == 0) hs = hash(aes, s);
    hs = shfl128(hs, 0);
    const uint32_t cs = (uint32_t)s.lo & 1u;
    if (lane < k) {
        const int j = lane;
        const ModInfo Pm = P.crt[j];
        const u128 G = gc::sub128(ldg128(P.garbler + et * (uint64_t)P.S + P.prefix[j] + col[j]), h[j]);
        const u128* ev = P.evaluator + (et * (uint64_t)k + j) * 3;
        const u128 E = gc::sub128(ldg128(ev + cs), hs);
        const u128 mini = ldg128(ev + 2);
        const int16_t row16 = (int16_t)((mini.lo >> (16 * cs)) & 0xffff);
        const int16_t t16 = (int16_t)(row16 - (int16_t)(hs.lo & 0xffff));
        int tt = (int)t16 % (int)Pm.q;
        if (tt < 0) tt += (int)Pm.q;
        const uint64_t off = plane_off(Pm, e, b, P.Bp);
        const uint8_t* xin = P.in[j] + off;
        uint8_t* xout = P.out[j] + off;
        uint32_t xw = 0, ow = 0;
        const uint32_t last = Pm.nc - 1;
        /* the input label's words again (they were read for the key in step
         * 1a), two ahead of their use */
        const uint32_t nw = nc4_of(Pm);
        auto xword = [&](uint32_t d4) { return d4 < nw ? *reinterpret_cast<const uint32_t*>(xin + (uint64_t)d4 * 4 * P.Bp) : 0u; };
        uint32_t xn0 = xword(0), xn1 = xword(1);
        gc::for_each_digit2(E, G, Pm, [&](uint32_t d, uint32_t de, uint32_t dg) {
            if ((d & 3) == 0) {
                xw = xn0;
                xn0 = xn1;
                xn1 = xword((d >> 2) + 2);
            }
            const uint32_t x = (xw >> (8 * (d & 3))) & 0xff;
            ow |= gc::mod_q(de + (uint32_t)tt * x + Pm.q - dg, Pm) << (8 * (d & 3));
            if ((d & 3) == 3 || d == last) {
                *reinterpret_cast<uint32_t*>(xout + (uint64_t)(d >> 2) * 4 * P.Bp) = ow;
                ow = 0;
            }
        });
    }
}

cudaError_t launch_sign(const SignParams& p, bool relu, cudaStream_t s) {
    if (p.n == 0) return cudaSuccess;
    const uint64_t total = p.n * p.batch;
    const unsigned grid = (unsigned)((total + SIGN_THREADS - 1) / SIGN_THREADS);
    uint32_t acc_digits = p.mrs[0].nc;
    for (int i = 1; i < p.n_mrs; ++i) acc_digits = max(acc_digits, p.cast[i].nc);
    /* k_sign: two digit sums per word; a modulus with a plan uses the stage-local pair layout */
    uint32_t acc_words = 0;
    for (int i = 0; i < p.n_mrs; ++i)
        acc_words = max(acc_words, (DASHB_FRAC && p.castf[i].nst) ? p.castf[i].words + 1 /* accumulate_t4 */ : (p.cast[i].nc + 1) / 2);
    const size_t smem = (size_t)acc_words * SIGN_THREADS * sizeof(uint32_t);
    /* latency mode: one warp per (element, inference) while that still leaves SMs idle otherwise */
    static const long wide_max = getenv("DASHB_WIDE_MAX") ? atol(getenv("DASHB_WIDE_MAX")) : 8192;
    /* with other lanes in flight the GPU is busy anyway: only tiny launches go wide */
    if ((long)total <= (p.lanes > 1 ? wide_max / 4 : wide_max) && p.k <= 32 && p.n_out <= 32 && p.rep <= 1) {
        const unsigned wgrid = (unsigned)((total + WIDE_WARPS - 1) / WIDE_WARPS);
        const size_t wsmem = (size_t)WIDE_WARPS * p.n_mrs * acc_digits * sizeof(uint32_t);
        if (wsmem <= 40 * 1024) {
            if (relu)
                k_sign_wide<true><<<wgrid, WIDE_WARPS * 32, wsmem, s>>>(p, acc_digits);
            else
                k_sign_wide<false><<<wgrid, WIDE_WARPS * 32, wsmem, s>>>(p, acc_digits);
            return cudaGetLastError();
        }
    }
    /* four threads per element while the launch could not fill the GPU anyway (148 SMs x 1024 threads):
     * a third of the latency for 1.4x the issue slots */
    static const long quad_max = getenv("DASHB_QUAD_MAX") ? atol(getenv("DASHB_QUAD_MAX")) : 75000;
    if ((long)total * (p.lanes > 1 ? p.lanes : 1) <= quad_max && p.k >= 4) {
        const unsigned qgrid = (unsigned)((total * 4 + SIGN_THREADS - 1) / SIGN_THREADS);
        if (relu)
            k_sign<true, 4><<<qgrid, SIGN_THREADS, smem, s>>>(p);
        else
            k_sign<false, 4><<<qgrid, SIGN_THREADS, smem, s>>>(p);
        return cudaGetLastError();
    }
    if (relu)
        k_sign<true, 1><<<grid, SIGN_THREADS, smem, s>>>(p);
    else
        k_sign<false, 1><<<grid, SIGN_THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

/* ------------------------------------------------------------ rescale */
/*
 * General-factor rescale with base extension, one thread per (element,
 * inference).  CPU-only in the reference ("Not yet implemented" on the GPU:
 * gadgets/base_extension_gadget.h:297-315; rescale_gadget.h:364); follows
 * RescaleGadget::cpu_evaluate (gadgets/rescale_gadget.h:244-362) and
 * BaseExtensionGadget::cpu_evaluate (gadgets/base_extension_gadget.h:
 * 197-295).  Each source label is hashed once for all the projections that
 * key on it.  With P.legacy the kernel stops after step 2.3; the legacy
 * variant's sign gadget (k_sign) and downshift (k_add_label) follow.
 *
 * Two rearrangements of the reference's arithmetic, results identical:
 *
 * (1) Column order.  The reference eliminates row by row (for every source
 *     residue, update all later residues), which needs the digits of all k
 *     residues at once.  Here one target residue at a time takes all the
 *     updates it will ever get, then is hashed (it becomes a source) or
 *     stored; a thread keeps one residue's digits plus the 16-byte hashes.
 *
 * (2) Lazy reduction.  A chain of updates w <- (w - p_i) * inv_i (mod q) is
 *     the linear form w0*A + sum_i p_i*B_i with A = prod inv_i and
 *     B_i = -prod_{i' >= i} inv_i' (host constants, RescaleParams::a1..b2).
 *     The digits of every projected label are multiplied by B_i and added
 *     to an unreduced accumulator (shared memory, [digit][thread]); one
 *     reduction mod q per digit at the end of the chain replaces one per
 *     update.  16-bit accumulators where (k+1)(q-1)^2 < 2^16, else 32-bit.
 */
constexpr int RS_THREADS = 128;
#ifndef DASHB_RS_MINB
#define DASHB_RS_MINB 8 /* CTAs per SM the register allocation aims at */
#endif

/* Accumulators of one thread, 32-bit words in shared memory, word-major
 * (acc[w * RS_THREADS]).  PACKED (the usual case, (k+1)(q-1)^2 < 2^16): two
 * digits per word, digit 2w in the low half and 2w+1 in the high half -- the
 * halves never meet, so one multiply-add updates both.  Otherwise one digit
 * per word. */

/* Two digits in the 16-bit halves of a word, each below 2q: subtract q from
 * the halves that reached it.  cq = (0x8000 - q) in both halves, so that bit 15
 * of (half + 0x8000 - q) says half >= q. */
__device__ __forceinline__ uint32_t pair_csub(uint32_t x, uint32_t q, uint32_t cq) {
    return x - (((x + cq) >> 15) & 0x00010001u) * q;
}
/* bytes 0,1 (HI = false) or 2,3 of a word of four digits as a pair of halves */
template <bool HI>
__device__ __forceinline__ uint32_t pair_of(uint32_t w) {
    return __byte_perm(w, 0, HI ? 0x4342 : 0x4140);
}

/* acc[d] = ((x_j[d] + upshift_j[d]) mod q) * a   (step 1, pre-multiplied) */
template <bool PACKED>
__device__ __forceinline__ void rs_init(uint32_t* acc, const uint8_t* plane, const uint8_t* up, const ModInfo& m, uint64_t e,
                                        uint32_t b, uint32_t Bp, uint32_t a) {
    const uint8_t* p = plane + plane_off(m, e, b, Bp);
    /* the label's words are requested two ahead of their use (see load_compress) */
    const uint32_t nw = nc4_of(m);
    auto word = [&](uint32_t d4) { return d4 < nw ? *reinterpret_cast<const uint32_t*>(p + (uint64_t)d4 * 4 * Bp) : 0u; };
    uint32_t n0 = word(0), n1 = word(1);
    const uint32_t cq = (0x8000u - m.q) * 0x10001u;
    for (uint32_t d4 = 0; d4 < nw; ++d4) {
        const uint32_t xw = n0;
        n0 = n1;
        n1 = word(d4 + 2);
        const uint32_t uw = __ldg(reinterpret_cast<const uint32_t*>(up) + d4);
        if (PACKED) {
            /* pairs of digits all the way: add, reduce and scale both halves at once (the padding digits of
             * the last word are zero in both labels and stay zero) */
            acc[(2 * d4) * RS_THREADS] = pair_csub(pair_of<false>(xw) + pair_of<false>(uw), m.q, cq) * a;
            acc[(2 * d4 + 1) * RS_THREADS] = pair_csub(pair_of<true>(xw) + pair_of<true>(uw), m.q, cq) * a;
        } else {
#pragma unroll
            for (uint32_t t = 0; t < 4; ++t) {
                uint32_t v = ((xw >> (8 * t)) & 0xff) + ((uw >> (8 * t)) & 0xff);
                v = (v >= m.q ? v - m.q : v) * a;
                if (4 * d4 + t < m.nc) acc[(4 * d4 + t) * RS_THREADS] = v;
            }
        }
    }
}

/* acc[d] = label[d] * a for a broadcast label (zero padded u8[4*nc4]) */
template <bool PACKED>
__device__ __forceinline__ void rs_set(uint32_t* acc, const uint8_t* label, const ModInfo& m, uint32_t a) {
    if (PACKED) {
        for (uint32_t w = 0; w < 2 * nc4_of(m); ++w)
            acc[w * RS_THREADS] = __ldg(label + 2 * w) * a | (__ldg(label + 2 * w + 1) * a) << 16;
    } else {
        for (uint32_t d = 0; d < m.nc; ++d) acc[d * RS_THREADS] = __ldg(label + d) * a;
    }
}

/* acc[d] += digit_d(pr) * bcoef by long division, least significant digit
 * first: the fallback for a modulus without a multiplicative plan.  Not
 * inlined, the modulus by value (in registers): one copy per kernel. */
template <bool PACKED>
__device__ __noinline__ void rs_accumulate(uint32_t* acc, u128 pr, const ModInfo m, uint32_t bcoef) {
    gc::for_each_digit(pr, m, [&](uint32_t d, uint32_t v) {
        if (PACKED)
            acc[(d >> 1) * RS_THREADS] += (v * bcoef) << (16 * (d & 1));
        else
            acc[d * RS_THREADS] += v * bcoef;
    });
}

/* The same with the digits taken by multiplication, most significant chunk
 * first (gc::for_each_chunk_msb; valid payloads).  Inlined: the plan is read
 * from the kernel-parameter bank, and the body is a fraction of the long
 * division's size.  PACKED: two digits cost two reciprocal multiplies, two
 * multiply-adds that also pack them (see accumulate_frac) and one
 * multiply-add into the shared-memory word. */
template <bool PACKED>
__device__ __forceinline__ void rs_accumulate_frac(uint32_t* acc, u128 pr, const ModInfo& m, const gc::FracInfo& fi,
                                                   uint32_t bcoef) {
    const uint32_t c1 = 65536u - m.q, c2 = m.q << 16;
    if (PACKED && fi.aligned) {
        /* word-aligned stages (make_fracinfo(aligned4)): every chunk starts on an even accumulator word, the
         * widths are known at compile time (gc::for_each_chunk_ladder) -- no parity case, no width dispatch */
        const uint32_t nc2 = 0u - c2, mq = m.mq;
        uint32_t d = m.nc;
        gc::for_each_chunk_ladder(pr, fi, [&](uint32_t len, uint32_t r) {
            d -= len;
            uint32_t* a = acc + (d >> 1) * RS_THREADS;
#pragma unroll 1
            for (; len >= 4; len -= 4) { /* two words per trip */
                const uint32_t q1 = gc::mulhi32(r, mq);
                const uint32_t q2 = gc::mulhi32(q1, mq);
                const uint32_t q3 = gc::mulhi32(q2, mq);
                const uint32_t q4 = gc::mulhi32(q3, mq);
                a[0] += (r + q1 * c1 + q2 * nc2) * bcoef;
                a[RS_THREADS] += (q2 + q3 * c1 + q4 * nc2) * bcoef;
                a += 2 * RS_THREADS;
                r = q4;
            }
            if (len & 2) {
                const uint32_t q1 = gc::mulhi32(r, mq);
                const uint32_t q2 = gc::mulhi32(q1, mq);
                a[0] += (r + q1 * c1 + q2 * nc2) * bcoef;
                a += RS_THREADS;
                r = q2;
            }
            if (len & 1) a[0] += r * bcoef; /* the chunk's last digit is what is left */
        });
        return;
    }
    gc::for_each_chunk_msb(pr, m, fi, [&](uint32_t d0, uint32_t len, uint32_t r) {
        uint32_t i = 0;
        if (PACKED) {
            uint32_t* a = acc + (d0 >> 1) * RS_THREADS;
            if (d0 & 1u) { /* the chunk starts in the high half of a word */
                const uint32_t q1 = gc::div_q(r, m);
                a[0] += ((r - q1 * m.q) * bcoef) << 16;
                a += RS_THREADS;
                r = q1;
                i = 1;
            }
#pragma unroll 1
            for (; i + 3 < len; i += 4) { /* two words per trip */
                const uint32_t q1 = gc::div_q(r, m);
                const uint32_t q2 = gc::div_q(q1, m);
                const uint32_t q3 = gc::div_q(q2, m);
                const uint32_t q4 = gc::div_q(q3, m);
                a[0] += (r + q1 * c1 - q2 * c2) * bcoef;
                a[RS_THREADS] += (q2 + q3 * c1 - q4 * c2) * bcoef;
                a += 2 * RS_THREADS;
                r = q4;
            }
            if (i + 1 < len) {
                const uint32_t q1 = gc::div_q(r, m);
                const uint32_t q2 = gc::div_q(q1, m);
                a[0] += (r + q1 * c1 - q2 * c2) * bcoef;
                a += RS_THREADS;
                r = q2;
                i += 2;
            }
            if (i < len) a[0] += r * bcoef; /* the chunk's last digit is what is left */
        } else {
            uint32_t* a = acc + d0 * RS_THREADS;
#pragma unroll 1
            for (; i + 1 < len; i += 2) {
                const uint32_t q1 = gc::div_q(r, m);
                const uint32_t q2 = gc::div_q(q1, m);
                a[0] += (r - q1 * m.q) * bcoef;
                a[RS_THREADS] += (q1 - q2 * m.q) * bcoef;
                a += 2 * RS_THREADS;
                r = q2;
            }
            if (i < len) a[0] += r * bcoef;
        }
    });
}

template <bool PACKED>
__device__ __forceinline__ uint32_t rs_field(const uint32_t* acc, uint32_t d) {
    if (!PACKED) return acc[d * RS_THREADS];
    const uint32_t w = acc[(d >> 1) * RS_THREADS];
    return (d & 1u) ? (w >> 16) : (w & 0xffffu);
}

/* reduce, compress and return the colour digit */
template <bool PACKED>
__device__ __forceinline__ u128 rs_finish_compress(const uint32_t* acc, const ModInfo& m, uint32_t& colour) {
    if (PACKED && m.cw) {
        /* four digits per trip, most significant word first: two accumulator words, four reductions, the
         * bytes of one label word, folded in by gc::WordCompressor (the top word's padding digits are zero) */
        gc::WordCompressor c(m);
        uint32_t v0 = 0;
        for (int d4 = (int)nc4_of(m) - 1; d4 >= 0; --d4) {
            const uint32_t w0 = acc[(2 * d4) * RS_THREADS], w1 = acc[(2 * d4 + 1) * RS_THREADS];
            v0 = gc::mod_q(w0 & 0xffffu, m);
            const uint32_t v1 = gc::mod_q(w0 >> 16, m), v2 = gc::mod_q(w1 & 0xffffu, m), v3 = gc::mod_q(w1 >> 16, m);
            c.push(v0 | (v1 << 8) | (v2 << 16) | (v3 << 24));
        }
        colour = v0;
        return c.value();
    }
    gc::Compressor c(m);
    uint32_t dg = 0;
    for (int d = (int)m.nc - 1; d >= 0; --d) {
        dg = gc::mod_q(rs_field<PACKED>(acc, (uint32_t)d), m);
        c.push(dg);
    }
    colour = dg;
    return c.value();
}

/* plane <- ((acc mod q) * mul - down) mod q; mul == 1 / down == nullptr skip
 * that step.  With next_a != 0 the accumulator restarts at (acc mod q) *
 * next_a for the next chain. */
template <bool PACKED>
__device__ __forceinline__ void rs_finish_store(uint32_t* acc, uint8_t* plane, const uint8_t* down, uint32_t mul, uint32_t next_a,
                                                const ModInfo& m, uint64_t e, uint32_t b, uint32_t Bp) {
    uint8_t* p = plane + plane_off(m, e, b, Bp);
    if (PACKED) {
        /* pairs of digits: the subtraction of the downshift label and the restart of the accumulator work on
         * both halves of a word at once; only the reductions mod q take the digits one by one.  Padding
         * digits are zero in the accumulator and in `down`, and come out zero. */
        const uint32_t cq = (0x8000u - m.q) * 0x10001u, qq = m.q * 0x10001u;
        for (uint32_t d4 = 0; d4 < nc4_of(m); ++d4) {
            const uint32_t w0 = acc[(2 * d4) * RS_THREADS], w1 = acc[(2 * d4 + 1) * RS_THREADS];
            uint32_t v0 = gc::mod_q(w0 & 0xffffu, m), v1 = gc::mod_q(w0 >> 16, m);
            uint32_t v2 = gc::mod_q(w1 & 0xffffu, m), v3 = gc::mod_q(w1 >> 16, m);
            if (next_a) {
                acc[(2 * d4) * RS_THREADS] = (v0 | (v1 << 16)) * next_a;
                acc[(2 * d4 + 1) * RS_THREADS] = (v2 | (v3 << 16)) * next_a;
            }
            if (mul != 1) {
                v0 = gc::mod_q(v0 * mul, m);
                v1 = gc::mod_q(v1 * mul, m);
                v2 = gc::mod_q(v2 * mul, m);
                v3 = gc::mod_q(v3 * mul, m);
            }
            uint32_t p0 = v0 | (v1 << 16), p1 = v2 | (v3 << 16);
            if (down) {
                const uint32_t dw = __ldg(reinterpret_cast<const uint32_t*>(down) + d4);
                p0 = pair_csub(p0 + qq - pair_of<false>(dw), m.q, cq);
                p1 = pair_csub(p1 + qq - pair_of<true>(dw), m.q, cq);
            }
            *reinterpret_cast<uint32_t*>(p + (uint64_t)d4 * 4 * Bp) = __byte_perm(p0, p1, 0x6420);
        }
        return;
    }
    for (uint32_t d4 = 0; d4 < nc4_of(m); ++d4) {
        const uint32_t dw = down ? __ldg(reinterpret_cast<const uint32_t*>(down) + d4) : 0u;
        uint32_t ow = 0;
#pragma unroll
        for (uint32_t t = 0; t < 4; ++t) {
            const uint32_t d = 4 * d4 + t;
            if (d >= m.nc) break;
            uint32_t v = gc::mod_q(acc[d * RS_THREADS], m);
            if (next_a) acc[d * RS_THREADS] = v * next_a;
            if (mul != 1) v = gc::mod_q(v * mul, m);
            if (down) {
                v = v + m.q - ((dw >> (8 * t)) & 0xff);
                v = v >= m.q ? v - m.q : v;
            }
            ow |= v << (8 * t);
        }
        *reinterpret_cast<uint32_t*>(p + (uint64_t)d4 * 4 * Bp) = ow;
    }
}

template <bool PACKED>
__device__ __forceinline__ void rescale_body(const RescaleParams& P, const AesSmem& aes, uint32_t* acc, uint64_t e, uint32_t b) {
    const int k = P.k;
    /* Steps 1-2 for the factors' own residues: the hash each factor's
     * trans-modulus projections key on (rescale_gadget.h:271-300).  A later
     * factor's residue has first been divided by the earlier factors. */
    u128 hf[MAXF];
    uint32_t fcol[MAXF];
#pragma unroll 1
    for (int i = 0; i < P.nf; ++i) {
        const int j = P.fidx[i];
        const ModInfo m = pin(P.crt[j]);
        rs_init<PACKED>(acc, P.in[j], P.up[j], m, e, b, P.Bp, P.a1f[i]);
#pragma unroll 1
        for (int i2 = 0; i2 < i; ++i2) {
            if (P.tm_off[i2][j] == ~0ull) continue;
            const u128 pr = gc::sub128(ldg128(P.tm + P.tm_off[i2][j] + e * (uint64_t)P.factor[i2] + fcol[i2]), hf[i2]);
            if (DASHB_FRAC && P.crtf[j].nst)
                rs_accumulate_frac<PACKED>(acc, pr, m, P.crtf[j], P.b1f[i][i2]);
            else
                rs_accumulate<PACKED>(acc, pr, m, P.b1f[i][i2]);
        }
        uint32_t colour;
        const u128 key = rs_finish_compress<PACKED>(acc, m, colour);
        hf[i] = hash(aes, key);
        fcol[i] = colour;
    }

    /* every residue in turn, in the base extension's swapped order
     * (base_extension_gadget.h:206-224: the factors' residues last) */
    u128 h[MAXK];
    uint32_t col[MAXK];
#pragma unroll 1
    for (int t = 0; t < k; ++t) {
        const int jt = P.legacy ? t : P.perm[t];
        const ModInfo m = pin(P.crt[jt]);
        int x = -1; /* index among the factors, -1 for the others */
        for (int i = 0; i < P.nf; ++i)
            if (P.fidx[i] == jt) x = i;
        if (x < 0) {
            /* steps 1-2: upshift, then divide by every factor.  The first
             * factor's ciphertext row is requested before the accumulator is
             * set up, the next one before the current one is consumed. */
            auto tm_row = [&](int i) {
                return i < P.nf && P.tm_off[i][jt] != ~0ull
                           ? ldg128(P.tm + P.tm_off[i][jt] + e * (uint64_t)P.factor[i] + fcol[i])
                           : gc::make_u128(0, 0);
            };
            u128 trow = tm_row(0);
            rs_init<PACKED>(acc, P.in[jt], P.up[jt], m, e, b, P.Bp, P.a1[jt]);
#pragma unroll 1
            for (int i = 0; i < P.nf; ++i) {
                const u128 cur = trow;
                trow = tm_row(i + 1);
                if (P.tm_off[i][jt] == ~0ull) continue;
                const u128 pr = gc::sub128(cur, hf[i]);
                if (DASHB_FRAC && P.crtf[jt].nst)
                    rs_accumulate_frac<PACKED>(acc, pr, m, P.crtf[jt], P.b1[i][jt]);
                else
                    rs_accumulate<PACKED>(acc, pr, m, P.b1[i][jt]);
            }
            /* step 4 for this residue (the base extension below only changes
             * the factors' residues, base_extension_gadget.h:253-294); the
             * accumulator restarts as the elimination's target */
            rs_finish_store<PACKED>(acc, P.out[jt], P.legacy ? nullptr : P.down[jt], 1, P.legacy ? 0u : (uint32_t)P.a2[t], m, e, b, P.Bp);
        } else if (P.legacy) {
            /* step 2.3: the factor's residue is the zero label */
            rs_set<PACKED>(acc, P.zero[jt], m, 1);
            rs_finish_store<PACKED>(acc, P.out[jt], nullptr, 1, 0, m, e, b, P.Bp);
        } else {
            rs_set<PACKED>(acc, P.zero[jt] + e * P.zstride[jt], m, P.a2[t]);
        }
        if (P.legacy) continue;
        /* step 3: garbled mixed-radix conversion, all eliminations of this target */
        const int ni = t < P.non_ext ? t : P.non_ext;
        auto be_row = [&](int i) {
            return i < ni ? ldg128(P.be + P.be_off[i][t] + e * (uint64_t)P.crt[P.perm[i]].q + col[i]) : gc::make_u128(0, 0);
        };
        u128 brow = be_row(0);
#pragma unroll 1
        for (int i = 0; i < ni; ++i) {
            const u128 pr = gc::sub128(brow, h[i]);
            brow = be_row(i + 1); /* the next row travels while this one is accumulated */
            if (DASHB_FRAC && P.crtf[jt].nst)
                rs_accumulate_frac<PACKED>(acc, pr, m, P.crtf[jt], P.b2[i][t]);
            else
                rs_accumulate<PACKED>(acc, pr, m, P.b2[i][t]);
        }
        if (x < 0) {
            uint32_t colour;
            const u128 key = rs_finish_compress<PACKED>(acc, m, colour);
            h[t] = hash(aes, key);
            col[t] = colour;
        } else {
            /* the extended residue is the last mixed-radix digit times -invv, then the downshift */
            rs_finish_store<PACKED>(acc, P.out[jt], P.down[jt], P.neg_invv[x], 0, m, e, b, P.Bp);
        }
    }
}

template <bool PACKED>
__global__ void __launch_bounds__(RS_THREADS, DASHB_RS_MINB) k_rescale(const __grid_constant__ RescaleParams P) {
    __shared__ AesSmem aes;
    extern __shared__ uint32_t s_rs[]; /* [2 * max nc4 (PACKED) or max nc][RS_THREADS] */
    aes_smem_load(aes);
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.n * P.batch) return;
    rescale_body<PACKED>(P, aes, s_rs + threadIdx.x, slice_elem(P.e0, P.seg, P.stride, idx / P.batch), (uint32_t)(idx % P.batch));
}

cudaError_t launch_rescale(const RescaleParams& p, cudaStream_t s) {
    if (p.n == 0) return cudaSuccess;
    const uint64_t total = p.n * p.batch;
    const unsigned grid = (unsigned)((total + RS_THREADS - 1) / RS_THREADS);
    uint32_t nc = 0;
    for (int j = 0; j < p.k; ++j) nc = max(nc, p.crt[j].nc);
    if (p.wide) {
        static bool opted_in[MAX_DEVICES];
        const int dev = initialised_device();
        if (dev >= 0 && dev < MAX_DEVICES && !opted_in[dev]) { /* 128 digits x 128 threads x 4 B */
            cudaError_t e = cudaFuncSetAttribute(k_rescale<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * RS_THREADS * 4);
            if (e != cudaSuccess) return e;
            opted_in[dev] = true;
        }
        k_rescale<false><<<grid, RS_THREADS, (size_t)nc * RS_THREADS * 4, s>>>(p);
    } else {
        k_rescale<true><<<grid, RS_THREADS, (size_t)((nc + 3) / 4 * 2) * RS_THREADS * 4, s>>>(p);
    }
    return cudaGetLastError();
}

/* AddLabel / SubLabel (misc/cuda_util.h:54-76): add or subtract one
 * broadcast label on every label of every plane. */
__global__ void k_add_label(const __grid_constant__ AddLabelParams P) {
    const int j = blockIdx.y;
    const ModInfo& m = P.crt[j];
    const uint32_t nc4 = (m.nc + 3) >> 2;
    const uint64_t first = P.e0 * nc4 * 4 * (uint64_t)P.Bp;
    const uint64_t total = first + P.n * nc4 * 4 * (uint64_t)P.Bp;
    for (uint64_t i = first + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t d = (uint32_t)((i / (4 * P.Bp)) % nc4) * 4 + (uint32_t)(i & 3);
        const uint32_t l = __ldg(P.label[j] + d); /* labels are zero-padded to 4*nc4 digits */
        uint32_t v = P.plane[j][i];
        v = P.subtract ? v + m.q - l : v + l;
        if (v >= m.q) v -= m.q;
        P.plane[j][i] = (uint8_t)v;
    }
}

cudaError_t launch_add_label(const AddLabelParams& p, cudaStream_t s) {
    if (p.n == 0) return cudaSuccess;
    uint64_t mx = 0;
    for (int j = 0; j < p.k; ++j) mx = max(mx, p.n * ((p.crt[j].nc + 3) & ~3u) * (uint64_t)p.Bp);
    uint64_t want = (mx + 255) / 256;
    unsigned gx = (unsigned)(want < 148 * 16 ? want : 148 * 16);
    dim3 grid(gx ? gx : 1, p.k);
    k_add_label<<<grid, 256, 0, s>>>(p);
    return cudaGetLastError();
}

/* ------------------------------------------------------ garble_inputs */
/* GarbledCircuitInterface::garble_inputs (garbling/
 * garbled_circuit_interface.h:325-343), batched and on the device:
 * label_j[e] = (x[e] mod p_j) * R_{p_j} + base_j[e], digit-wise mod p_j
 * (util::crt_reduce gives the non-negative residue, misc/util.h:143-156).
 * One thread per (input, inference) and modulus; writes the plane words. */
__global__ void __launch_bounds__(128) k_garble_inputs(const __grid_constant__ GarbleInParams P) {
    const int j = blockIdx.y;
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.n * P.batch) return;
    const uint32_t b = (uint32_t)(idx % P.batch);
    const uint64_t e = idx / P.batch;
    const ModInfo m = pin(P.crt[j]);
    const int64_t q = (int64_t)m.q;
    int64_t r64 = P.x[(uint64_t)b * P.n + e] % q;
    if (r64 < 0) r64 += q;
    const uint32_t r = (uint32_t)r64;
    const uint32_t nc4 = nc4_of(m);
    const uint32_t* base = reinterpret_cast<const uint32_t*>(P.base[j]) + e * nc4;
    const uint32_t* off = reinterpret_cast<const uint32_t*>(P.off[j]);
    uint8_t* out = P.out[j] + plane_off(m, e, b, P.Bp);
    for (uint32_t d4 = 0; d4 < nc4; ++d4) {
        const uint32_t bw = __ldg(base + d4), ow = __ldg(off + d4);
        uint32_t w = 0;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const uint32_t d = 4 * d4 + t;
            const uint32_t v = gc::mod_q(r * ((ow >> (8 * t)) & 0xff) + ((bw >> (8 * t)) & 0xff), m);
            w |= (d < m.nc ? v : 0u) << (8 * t);
        }
        *reinterpret_cast<uint32_t*>(out + (uint64_t)d4 * 4 * P.Bp) = w;
    }
}

cudaError_t launch_garble_inputs(const GarbleInParams& p, cudaStream_t s) {
    const uint64_t total = p.n * p.batch;
    dim3 grid((unsigned)((total + 127) / 128), p.k);
    k_garble_inputs<<<grid, 128, 0, s>>>(p);
    return cudaGetLastError();
}

/* ------------------------------------------------------------- decode */
/* decode_outputs (garbling/garbled_circuit_interface.h:422-467): hash each
 * output label, find the residue whose decoding hash matches, CRT-combine,
 * map to the signed range.  One thread per (output, inference). */
__global__ void __launch_bounds__(128) k_decode(const __grid_constant__ DecodeParams P) {
    __shared__ AesSmem aes;
    aes_smem_load(aes);
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.n * P.batch) return;
    const uint32_t b = (uint32_t)(idx % P.batch);
    const uint64_t e = idx / P.batch;
    unsigned __int128 sum = 0;
    bool good = true;
    for (int j = 0; j < P.k; ++j) {
        uint32_t colour;
        const u128 hh = hash(aes, load_compress(P.in[j], P.crt[j], e, b, P.Bp, colour));
        int found = -1;
        for (uint32_t r = 0; r < P.crt[j].q; ++r) {
            const u128 dref = ldg128(P.dec[j] + (uint64_t)r * P.n + e);
            if (dref.lo == hh.lo && dref.hi == hh.hi) found = (int)r;
        }
        if (found < 0) {
            good = false;
            found = 0;
        }
        /* util::chinese_remainder (misc/util.h:116-141) term by term */
        const unsigned __int128 prod = (unsigned __int128)(uint64_t)P.prod;
        sum += (((unsigned __int128)(uint64_t)found * (uint64_t)P.inv[j]) % prod * (uint64_t)P.pdiv[j]) % prod;
    }
    int64_t v = (int64_t)(uint64_t)(sum % (uint64_t)P.prod);
    if (v >= P.crt_modulus / 2) v -= P.crt_modulus;
    P.values[(uint64_t)b * P.n + e] = v;
    if (!good) P.ok[b] = 0;
}

cudaError_t launch_decode(const DecodeParams& p, cudaStream_t s) {
    const uint64_t total = p.n * p.batch;
    k_decode<<<(unsigned)((total + 127) / 128), 128, 0, s>>>(p);
    return cudaGetLastError();
}

}  // namespace dashb

#include "garble_kernels.cuh"
#include "shard_kernels.cuh"
