// AES-only kernel peak on B200 ("AES blocks/s divided by
// a measured AES-only kernel peak"), and a test bed for formulations of the
// T-table round.  Every thread runs a dependent chain of fixed-key AES-128
// encryptions, the way the gadget kernels hash labels; a host AES checks the
// chain of thread 0.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o aes_probe tools/aes_probe.cu
//   ./aes_probe            -> one line per variant: blocks/s
//
// Variants:
//   0  shift+mask index, four tables (what dash_b200/csrc/gc_common.cuh does)
//   1  byte extraction with PRMT, scale+base as an IMAD (FMA pipe)
//   2  variant 1 + round keys as one 128-bit constant load per round
//   3  variant 2 with the nine rounds unrolled by three
//   4  ONE table, replicated 32 times so that lane l only ever touches bank l
//      (entry i of the lane's copy at word 32*i + l: no bank conflicts; 32 KB
//      per CTA); T1..T3 are byte rotations (PRMT) of the T0 entry and the
//      last round takes the S-box byte out of the same entry
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

static int g_reps = 10, g_warm = 3;
static uint8_t h_sbox[256];
static uint32_t h_te[4][256];
static uint32_t h_rk[44];

static uint8_t xt(uint8_t x) { return (uint8_t)((x << 1) ^ ((x & 0x80) ? 0x1b : 0)); }

static void build_tables() {
    // S-box from the multiplicative inverse in GF(2^8) and the affine map
    uint8_t p = 1, q = 1;
    do {
        p = p ^ (uint8_t)(p << 1) ^ ((p & 0x80) ? 0x1b : 0);
        q ^= q << 1;
        q ^= q << 2;
        q ^= q << 4;
        if (q & 0x80) q ^= 0x09;
        uint8_t x = q ^ (uint8_t)((q << 1) | (q >> 7)) ^ (uint8_t)((q << 2) | (q >> 6)) ^ (uint8_t)((q << 3) | (q >> 5)) ^
                    (uint8_t)((q << 4) | (q >> 4));
        h_sbox[p] = x ^ 0x63;
    } while (p != 1);
    h_sbox[0] = 0x63;
    for (int i = 0; i < 256; ++i) {
        uint8_t s = h_sbox[i], s2 = xt(s), s3 = s2 ^ s;
        uint32_t t = ((uint32_t)s2 << 24) | ((uint32_t)s << 16) | ((uint32_t)s << 8) | s3;
        h_te[0][i] = t;
        h_te[1][i] = (t >> 8) | (t << 24);
        h_te[2][i] = (t >> 16) | (t << 16);
        h_te[3][i] = (t >> 24) | (t << 8);
    }
    static const uint8_t rcon[10] = {1, 2, 4, 8, 16, 32, 64, 128, 0x1b, 0x36};
    for (int i = 0; i < 4; ++i) h_rk[i] = ((4u * i) << 24) | ((4u * i + 1) << 16) | ((4u * i + 2) << 8) | (4u * i + 3);
    for (int i = 4; i < 44; ++i) {
        uint32_t t = h_rk[i - 1];
        if (i % 4 == 0) {
            t = (t << 8) | (t >> 24);
            t = ((uint32_t)h_sbox[t >> 24] << 24) | ((uint32_t)h_sbox[(t >> 16) & 255] << 16) |
                ((uint32_t)h_sbox[(t >> 8) & 255] << 8) | h_sbox[t & 255];
            t ^= (uint32_t)rcon[i / 4 - 1] << 24;
        }
        h_rk[i] = h_rk[i - 4] ^ t;
    }
}

static void host_aes(uint32_t s[4]) {
    for (int i = 0; i < 4; ++i) s[i] ^= h_rk[i];
    for (int r = 1; r < 10; ++r) {
        uint32_t t[4];
        for (int c = 0; c < 4; ++c)
            t[c] = h_te[0][s[c] >> 24] ^ h_te[1][(s[(c + 1) & 3] >> 16) & 255] ^ h_te[2][(s[(c + 2) & 3] >> 8) & 255] ^
                   h_te[3][s[(c + 3) & 3] & 255] ^ h_rk[4 * r + c];
        for (int c = 0; c < 4; ++c) s[c] = t[c];
    }
    uint32_t o[4];
    for (int c = 0; c < 4; ++c)
        o[c] = ((uint32_t)h_sbox[s[c] >> 24] << 24) ^ ((uint32_t)h_sbox[(s[(c + 1) & 3] >> 16) & 255] << 16) ^
               ((uint32_t)h_sbox[(s[(c + 2) & 3] >> 8) & 255] << 8) ^ h_sbox[s[(c + 3) & 3] & 255] ^ h_rk[40 + c];
    for (int c = 0; c < 4; ++c) s[c] = o[c];
}

__constant__ __align__(16) uint32_t c_rk[44];
__device__ uint32_t g_te0[256];
__device__ uint8_t g_sbox[256];

struct Smem {
    uint32_t te[4][256];
    uint8_t sbox[256];
};

__device__ __forceinline__ uint32_t byte_of(uint32_t s, int i) {
    return i == 0 ? (s & 0xffu) : __byte_perm(s, 0, 0x4440 + i);
}

template <int V>
__device__ __forceinline__ uint32_t lut(const Smem& a, int tab, uint32_t s, int byte, uint32_t four) {
    if (V == 0) {
        const uint32_t off = byte == 0 ? (s << 2) & 0x3fcu : (s >> (8 * byte - 2)) & 0x3fcu;
        return *reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(a.te[tab]) + off);
    } else {
        // shared-window address of te[tab] + 4 * byte, the multiply-add on the FMA pipe
        const uint32_t base = (uint32_t)__cvta_generic_to_shared(a.te[tab]);
        uint32_t addr, v;
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(addr) : "r"(byte_of(s, byte)), "r"(four), "r"(base));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
        return v;
    }
}

template <int V>
__device__ __forceinline__ void round_fn(const Smem& a, uint32_t& s0, uint32_t& s1, uint32_t& s2, uint32_t& s3, int r,
                                         uint32_t four) {
    uint32_t k0, k1, k2, k3;
    if (V >= 2) {
        const uint4 k4 = reinterpret_cast<const uint4*>(c_rk)[r];
        k0 = k4.x, k1 = k4.y, k2 = k4.z, k3 = k4.w;
    } else {
        k0 = c_rk[4 * r], k1 = c_rk[4 * r + 1], k2 = c_rk[4 * r + 2], k3 = c_rk[4 * r + 3];
    }
    const uint32_t t0 = lut<V>(a, 0, s0, 3, four) ^ lut<V>(a, 1, s1, 2, four) ^ lut<V>(a, 2, s2, 1, four) ^ lut<V>(a, 3, s3, 0, four) ^ k0;
    const uint32_t t1 = lut<V>(a, 0, s1, 3, four) ^ lut<V>(a, 1, s2, 2, four) ^ lut<V>(a, 2, s3, 1, four) ^ lut<V>(a, 3, s0, 0, four) ^ k1;
    const uint32_t t2 = lut<V>(a, 0, s2, 3, four) ^ lut<V>(a, 1, s3, 2, four) ^ lut<V>(a, 2, s0, 1, four) ^ lut<V>(a, 3, s1, 0, four) ^ k2;
    const uint32_t t3 = lut<V>(a, 0, s3, 3, four) ^ lut<V>(a, 1, s0, 2, four) ^ lut<V>(a, 2, s1, 1, four) ^ lut<V>(a, 3, s2, 0, four) ^ k3;
    s0 = t0, s1 = t1, s2 = t2, s3 = t3;
}

template <int V>
__device__ __noinline__ uint4 aes(const Smem& a, uint4 in, uint32_t four) {
    uint32_t s0 = in.x ^ c_rk[0], s1 = in.y ^ c_rk[1], s2 = in.z ^ c_rk[2], s3 = in.w ^ c_rk[3];
    if (V == 3) {
#pragma unroll 1
        for (int r = 1; r < 10; r += 3) {
            round_fn<V>(a, s0, s1, s2, s3, r, four);
            round_fn<V>(a, s0, s1, s2, s3, r + 1, four);
            round_fn<V>(a, s0, s1, s2, s3, r + 2, four);
        }
    } else {
#pragma unroll 1
        for (int r = 1; r < 10; ++r) round_fn<V>(a, s0, s1, s2, s3, r, four);
    }
    const uint8_t* sb = a.sbox;
    const uint32_t o0 = ((uint32_t)sb[s0 >> 24] << 24) ^ ((uint32_t)sb[(s1 >> 16) & 255] << 16) ^ ((uint32_t)sb[(s2 >> 8) & 255] << 8) ^ sb[s3 & 255] ^ c_rk[40];
    const uint32_t o1 = ((uint32_t)sb[s1 >> 24] << 24) ^ ((uint32_t)sb[(s2 >> 16) & 255] << 16) ^ ((uint32_t)sb[(s3 >> 8) & 255] << 8) ^ sb[s0 & 255] ^ c_rk[41];
    const uint32_t o2 = ((uint32_t)sb[s2 >> 24] << 24) ^ ((uint32_t)sb[(s3 >> 16) & 255] << 16) ^ ((uint32_t)sb[(s0 >> 8) & 255] << 8) ^ sb[s1 & 255] ^ c_rk[42];
    const uint32_t o3 = ((uint32_t)sb[s3 >> 24] << 24) ^ ((uint32_t)sb[(s0 >> 16) & 255] << 16) ^ ((uint32_t)sb[(s1 >> 8) & 255] << 8) ^ sb[s2 & 255] ^ c_rk[43];
    return make_uint4(o0, o1, o2, o3);
}

template <int V>
__global__ void __launch_bounds__(128) k_chain(uint4* out, int chain, uint32_t four) {
    __shared__ Smem a;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        const uint32_t t = g_te0[i];
        a.te[0][i] = t;
        a.te[1][i] = (t >> 8) | (t << 24);
        a.te[2][i] = (t >> 16) | (t << 16);
        a.te[3][i] = (t >> 24) | (t << 8);
        a.sbox[i] = g_sbox[i];
    }
    __syncthreads();
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    uint4 s = make_uint4(id, id * 2654435761u, ~id, 0x9e3779b9u ^ id);
#pragma unroll 1
    for (int i = 0; i < chain; ++i) s = aes<V>(a, s, four);
    out[id] = s;
}

#define CK(x)                                                                   \
    do {                                                                        \
        cudaError_t e_ = (x);                                                   \
        if (e_ != cudaSuccess) {                                                \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));            \
            return 1;                                                           \
        }                                                                       \
    } while (0)

__device__ __forceinline__ uint32_t lut_rep(uint32_t lane_base, uint32_t s, int byte, uint32_t stride) {
    uint32_t addr, v;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(addr) : "r"(byte_of(s, byte)), "r"(stride), "r"(lane_base));
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ uint32_t col_rep(uint32_t lb, uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t k, uint32_t st) {
    /* T0[a3] ^ ror8(T0[b2]) ^ ror16(T0[c1]) ^ ror24(T0[d0]) ^ k */
    const uint32_t x0 = lut_rep(lb, a, 3, st);
    const uint32_t x1 = __byte_perm(lut_rep(lb, b, 2, st), 0, 0x0321);
    const uint32_t x2 = __byte_perm(lut_rep(lb, c, 1, st), 0, 0x1032);
    const uint32_t x3 = __byte_perm(lut_rep(lb, d, 0, st), 0, 0x2103);
    return x0 ^ x1 ^ x2 ^ x3 ^ k;
}

/* S-box byte (byte 1 of the T0 entry) moved to byte position `pos` */
__device__ __forceinline__ uint32_t sb_rep(uint32_t lb, uint32_t s, int byte, int pos, uint32_t st) {
    const uint32_t t = lut_rep(lb, s, byte, st);
    return __byte_perm(t, 0, pos == 3 ? 0x1444 : pos == 2 ? 0x4144 : pos == 1 ? 0x4414 : 0x4441);
}

__device__ __noinline__ uint4 aes_rep(uint32_t lb, uint4 in, uint32_t st) {
    uint32_t s0 = in.x ^ c_rk[0], s1 = in.y ^ c_rk[1], s2 = in.z ^ c_rk[2], s3 = in.w ^ c_rk[3];
#pragma unroll 1
    for (int r = 1; r < 10; ++r) {
        const uint4 k4 = reinterpret_cast<const uint4*>(c_rk)[r];
        const uint32_t t0 = col_rep(lb, s0, s1, s2, s3, k4.x, st);
        const uint32_t t1 = col_rep(lb, s1, s2, s3, s0, k4.y, st);
        const uint32_t t2 = col_rep(lb, s2, s3, s0, s1, k4.z, st);
        const uint32_t t3 = col_rep(lb, s3, s0, s1, s2, k4.w, st);
        s0 = t0, s1 = t1, s2 = t2, s3 = t3;
    }
    const uint32_t o0 = sb_rep(lb, s0, 3, 3, st) ^ sb_rep(lb, s1, 2, 2, st) ^ sb_rep(lb, s2, 1, 1, st) ^ sb_rep(lb, s3, 0, 0, st) ^ c_rk[40];
    const uint32_t o1 = sb_rep(lb, s1, 3, 3, st) ^ sb_rep(lb, s2, 2, 2, st) ^ sb_rep(lb, s3, 1, 1, st) ^ sb_rep(lb, s0, 0, 0, st) ^ c_rk[41];
    const uint32_t o2 = sb_rep(lb, s2, 3, 3, st) ^ sb_rep(lb, s3, 2, 2, st) ^ sb_rep(lb, s0, 1, 1, st) ^ sb_rep(lb, s1, 0, 0, st) ^ c_rk[42];
    const uint32_t o3 = sb_rep(lb, s3, 3, 3, st) ^ sb_rep(lb, s0, 2, 2, st) ^ sb_rep(lb, s1, 1, 1, st) ^ sb_rep(lb, s2, 0, 0, st) ^ c_rk[43];
    return make_uint4(o0, o1, o2, o3);
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS) k_chain_rep(uint4* out, int chain, uint32_t stride) {
    extern __shared__ uint32_t rep[]; /* [256][32] */
    for (int i = threadIdx.x; i < 256 * 32; i += blockDim.x) rep[i] = g_te0[i >> 5];
    __syncthreads();
    const uint32_t lb = (uint32_t)__cvta_generic_to_shared(rep) + (threadIdx.x & 31) * 4;
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    uint4 s = make_uint4(id, id * 2654435761u, ~id, 0x9e3779b9u ^ id);
#pragma unroll 1
    for (int i = 0; i < chain; ++i) s = aes_rep(lb, s, stride);
    out[id] = s;
}

template <int THREADS>
static int run_rep(int total_threads, int chain, uint4* d_out) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    const int ctas = total_threads / THREADS;
    CK(cudaFuncSetAttribute(k_chain_rep<THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768));
    for (int w = 0; w < g_warm; ++w) k_chain_rep<THREADS><<<ctas, THREADS, 32768>>>(d_out, chain, 128);
    CK(cudaDeviceSynchronize());
    const int reps = g_reps;
    CK(cudaEventRecord(e0));
    for (int r = 0; r < reps; ++r) k_chain_rep<THREADS><<<ctas, THREADS, 32768>>>(d_out, chain, 128);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    uint4 got;
    CK(cudaMemcpy(&got, d_out, sizeof(got), cudaMemcpyDeviceToHost));
    uint32_t s[4] = {0u, 0u, ~0u, 0x9e3779b9u};
    for (int i = 0; i < chain; ++i) host_aes(s);
    const bool ok = got.x == s[0] && got.y == s[1] && got.z == s[2] && got.w == s[3];
    printf("{\"variant\": 4, \"threads_per_cta\": %d, \"ctas\": %d, \"chain\": %d, \"ms_per_launch\": %.4f, \"aes_blocks_per_s\": %.4g, \"matches_host_aes\": %s}\n",
           THREADS, ctas, chain, ms / reps, (double)ctas * THREADS * chain * reps / (ms / 1e3), ok ? "true" : "false");
    return ok ? 0 : 2;
}

template <int V>
static int run(int ctas, int chain, uint4* d_out, double* rate) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int w = 0; w < g_warm; ++w) k_chain<V><<<ctas, 128>>>(d_out, chain, 4);
    CK(cudaDeviceSynchronize());
    const int reps = g_reps;
    CK(cudaEventRecord(e0));
    for (int r = 0; r < reps; ++r) k_chain<V><<<ctas, 128>>>(d_out, chain, 4);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    *rate = (double)ctas * 128 * chain * reps / (ms / 1e3);
    uint4 got;
    CK(cudaMemcpy(&got, d_out, sizeof(got), cudaMemcpyDeviceToHost));
    uint32_t s[4] = {0u, 0u, ~0u, 0x9e3779b9u};
    for (int i = 0; i < chain; ++i) host_aes(s);
    const bool ok = got.x == s[0] && got.y == s[1] && got.z == s[2] && got.w == s[3];
    printf("{\"variant\": %d, \"ctas\": %d, \"chain\": %d, \"ms_per_launch\": %.4f, \"aes_blocks_per_s\": %.4g, \"matches_host_aes\": %s}\n",
           V, ctas, chain, ms / reps, *rate, ok ? "true" : "false");
    return ok ? 0 : 2;
}

int main(int argc, char** argv) {
    build_tables();
    const int chain = argc > 1 ? atoi(argv[1]) : 88;
    if (argc > 2) g_reps = atoi(argv[2]), g_warm = 0; /* for a profiler: no warm-up, few launches */
    int sms = 148;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    sms = prop.multiProcessorCount;
    const int ctas = sms * 16 * 4; /* four full waves at 16 CTAs of 128 threads per SM */
    CK(cudaMemcpyToSymbol(c_rk, h_rk, sizeof(h_rk)));
    CK(cudaMemcpyToSymbol(g_te0, h_te[0], sizeof(h_te[0])));
    CK(cudaMemcpyToSymbol(g_sbox, h_sbox, sizeof(h_sbox)));
    uint4* d_out = nullptr;
    CK(cudaMalloc(&d_out, (size_t)ctas * 128 * sizeof(uint4)));
    double r;
    int rc = 0;
    rc |= run<0>(ctas, chain, d_out, &r);
    rc |= run<1>(ctas, chain, d_out, &r);
    rc |= run<2>(ctas, chain, d_out, &r);
    rc |= run<3>(ctas, chain, d_out, &r);
    rc |= run_rep<256>(ctas * 128, chain, d_out);
    rc |= run_rep<512>(ctas * 128, chain, d_out);
    rc |= run_rep<1024>(ctas * 128, chain, d_out);
    cudaFree(d_out);
    return rc;
}
