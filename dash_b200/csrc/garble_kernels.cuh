/*
 * garble_kernels.cuh -- the offline phase on the GPU: one
 * generic kernel that garbles projection gates, included at the end of
 * kernels.cu so that it shares the hash tables and device helpers.
 *
 * The reference garbles gate by gate on the host: ProjectionGate::garble
 * (garbling/gates/projection_gate.h:62-87) draws a random output base label
 * and, for every input value v, stores
 *     row[colour(K_v)] = compress(base_out + f(v) * offset) + H(compress(K_v)),
 *     K_v = base_in + v * R_in,
 * where base_in is whatever label arithmetic the surrounding gadget did on
 * earlier gates' base labels (sign_gadget.h:420-581, rescale_gadget.h:114-243,
 * base_extension_gadget.h:63-196, mixed_mod_half_gate.h:51-89).
 *
 * Here every fresh base label is PRF_seed(layer, tag, element) (AES-128 under
 * a key derived from the garbling seed), so ANY gate of a layer can be
 * garbled independently of every other: its input label is a short linear
 * combination of (a) the layer's input base labels, (b) circuit-wide
 * constant labels and (c) PRF labels recomputed on the spot.  The host turns
 * a gadget into a list of gate descriptors (GateDesc, garbler.cu), the kernel
 * runs one thread per (descriptor, element); a warp holds 32 elements of the
 * same descriptor, so its lanes loop over the same number of rows.
 */

namespace dashb {

constexpr int GB_THREADS = 64;

__constant__ __align__(16) uint32_t c_rk_prf[44];

cudaError_t garbler_set_prf_key(const uint32_t rk[44]) { return cudaMemcpyToSymbol(c_rk_prf, rk, 44 * sizeof(uint32_t)); }

/* AES-128 under the garbling key: the PRF behind every random label. */
__device__ __noinline__ u128 prf(const AesSmem& a, u128 x) {
    const uint32_t four = c_four;
    uint32_t s0 = gc::bswap32((uint32_t)x.lo) ^ c_rk_prf[0];
    uint32_t s1 = gc::bswap32((uint32_t)(x.lo >> 32)) ^ c_rk_prf[1];
    uint32_t s2 = gc::bswap32((uint32_t)x.hi) ^ c_rk_prf[2];
    uint32_t s3 = gc::bswap32((uint32_t)(x.hi >> 32)) ^ c_rk_prf[3];
#pragma unroll 1
    for (int r = 1; r < 10; ++r) {
        const uint4 k4 = reinterpret_cast<const uint4*>(c_rk_prf)[r];
        const uint32_t t0 = te_at<0, 3>(a, s0, four) ^ te_at<1, 2>(a, s1, four) ^ te_at<2, 1>(a, s2, four) ^ te_at<3, 0>(a, s3, four) ^ k4.x;
        const uint32_t t1 = te_at<0, 3>(a, s1, four) ^ te_at<1, 2>(a, s2, four) ^ te_at<2, 1>(a, s3, four) ^ te_at<3, 0>(a, s0, four) ^ k4.y;
        const uint32_t t2 = te_at<0, 3>(a, s2, four) ^ te_at<1, 2>(a, s3, four) ^ te_at<2, 1>(a, s0, four) ^ te_at<3, 0>(a, s1, four) ^ k4.z;
        const uint32_t t3 = te_at<0, 3>(a, s3, four) ^ te_at<1, 2>(a, s0, four) ^ te_at<2, 1>(a, s1, four) ^ te_at<3, 0>(a, s2, four) ^ k4.w;
        s0 = t0, s1 = t1, s2 = t2, s3 = t3;
    }
    const uint32_t o0 = (sbox_at<3>(a, s0) << 24) ^ (sbox_at<2>(a, s1) << 16) ^ (sbox_at<1>(a, s2) << 8) ^ sbox_at<0>(a, s3) ^ c_rk_prf[40];
    const uint32_t o1 = (sbox_at<3>(a, s1) << 24) ^ (sbox_at<2>(a, s2) << 16) ^ (sbox_at<1>(a, s3) << 8) ^ sbox_at<0>(a, s0) ^ c_rk_prf[41];
    const uint32_t o2 = (sbox_at<3>(a, s2) << 24) ^ (sbox_at<2>(a, s3) << 16) ^ (sbox_at<1>(a, s0) << 8) ^ sbox_at<0>(a, s1) ^ c_rk_prf[42];
    const uint32_t o3 = (sbox_at<3>(a, s3) << 24) ^ (sbox_at<2>(a, s0) << 16) ^ (sbox_at<1>(a, s1) << 8) ^ sbox_at<0>(a, s2) ^ c_rk_prf[43];
    u128 out;
    out.lo = (uint64_t)o0 | ((uint64_t)o1 << 32);
    out.hi = (uint64_t)o2 | ((uint64_t)o3 << 32);
    return out;
}

/* acc[d] = (acc[d] + coef * digit_d(x)) mod q for the nc digits of the
 * 128-bit value x -- a PRF output read as a label (its low nc-1 digits are
 * uniform, the top one as uniform as 2^128 / q^(nc-1) allows). */
__device__ __noinline__ void gb_add_value(uint16_t* acc, u128 x, const ModInfo m, uint32_t coef) {
    gc::for_each_digit(x, m, [&](uint32_t d, uint32_t v) {
        acc[d * GB_THREADS] = (uint16_t)gc::mod_q(acc[d * GB_THREADS] + coef * v, m);
    });
}

/* same for a label stored as u8 digits with `stride` bytes between them */
__device__ __forceinline__ void gb_add_digits(uint16_t* acc, const uint8_t* lab, const ModInfo& m, uint32_t coef) {
    for (uint32_t d = 0; d < m.nc; ++d)
        acc[d * GB_THREADS] = (uint16_t)gc::mod_q(acc[d * GB_THREADS] + coef * __ldg(lab + d), m);
}

/* label of plane j (garbler layout: u8 [element][4*nc4]) */
__device__ __forceinline__ const uint8_t* gb_plane_label(const GarbleParams& P, uint32_t j, uint64_t e) {
    return P.in[j] + e * (uint64_t)(((P.in_nc[j] + 3) >> 2) << 2);
}

__device__ __forceinline__ u128 gb_counter(const GarbleParams& P, uint64_t tag, uint64_t e) {
    return gc::make_u128(e, (P.layer_tag << 32) | tag);
}

/* acc = sum of the terms, as a label of modulus m */
__device__ __forceinline__ void gb_combine(const GarbleParams& P, const AesSmem& aes, uint16_t* acc, const GateTerm* terms,
                                           uint32_t n_terms, const ModInfo& m, uint64_t e) {
    for (uint32_t d = 0; d < m.nc; ++d) acc[d * GB_THREADS] = 0;
#pragma unroll 1
    for (uint32_t t = 0; t < n_terms; ++t) {
        const GateTerm T = terms[t];
        if (T.kind == GB_TERM_PLANE)
            gb_add_digits(acc, gb_plane_label(P, (uint32_t)T.arg, e), m, T.coef);
        else if (T.kind == GB_TERM_CONST)
            gb_add_digits(acc, P.labels + T.arg, m, T.coef);
        else
            gb_add_value(acc, prf(aes, gb_counter(P, T.arg, e)), m, T.coef);
    }
}

__device__ __forceinline__ uint32_t gb_function(const GateDesc& G, uint32_t v, uint32_t r, uint32_t q_out) {
    switch (G.f) {
        case GB_F_ID: return v;
        case GB_F_LOOKUP: return __ldg(reinterpret_cast<const uint8_t*>(G.flut) + (uint64_t)v * G.flut_stride);
        case GB_F_DIV: return v / (uint32_t)G.fa;
        case GB_F_SIGN: { /* sign_functionality (sign_gadget.h:34-38): fa = m_0 / 2, fb = upper, fc = lower */
            const int x = v < (uint32_t)G.fa ? G.fb : G.fc;
            const int y = x % (int)q_out;
            return (uint32_t)(y < 0 ? y + (int)q_out : y);
        }
        case GB_F_MULR: return (v * r) % q_out;                               /* garbler half gate: x * r */
        case GB_F_NEGSUM: return (q_out - (v + r) % q_out) % q_out;           /* evaluator half gate: -(y + r) */
        case GB_F_ADDR: return (v + r) % (uint32_t)G.fa;                      /* mini gate: (y + r) mod p */
        default: return 0;
    }
}

__global__ void __launch_bounds__(GB_THREADS) k_garble_gates(const __grid_constant__ GarbleParams P) {
    __shared__ AesSmem aes;
    extern __shared__ uint16_t s_gb[]; /* key digits [nc_max][T], then payload base digits [nc_max][T] */
    aes_smem_load(aes);
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= P.n) return;
    const GateDesc& G = P.desc[blockIdx.y];
    uint16_t* key = s_gb + threadIdx.x;
    uint16_t* base = s_gb + (size_t)P.nc_max * GB_THREADS + threadIdx.x;

    /* out label (the gate's fresh base, or a plain label combination) */
    ModInfo mo;
    if (G.q_out) {
        mo = P.mods[G.q_out];
        gb_combine(P, aes, base, G.out, G.n_out, mo, e);
        if (G.out_plane >= 0) {
            uint8_t* dst = P.out[G.out_plane] + e * (uint64_t)(((mo.nc + 3) >> 2) << 2);
            for (uint32_t d = 0; d < (((mo.nc + 3) >> 2) << 2); ++d) dst[d] = d < mo.nc ? (uint8_t)base[d * GB_THREADS] : 0;
        }
    }
    if (!G.q_in) return;

    const ModInfo mi = P.mods[G.q_in];
    gb_combine(P, aes, key, G.in, G.n_in, mi, e);
    const uint8_t* r_in = P.labels + P.r_off[G.q_in];
    /* payload offset label: R_{q_out}, or (evaluator half gate) the label of an input plane */
    const uint8_t* off = nullptr;
    if (G.q_out) off = G.off_plane == GB_NONE ? P.labels + P.r_off[G.q_out] : gb_plane_label(P, G.off_plane, e);
    const uint32_t r = G.r_plane == GB_NONE ? 0u : (uint32_t)__ldg(gb_plane_label(P, G.r_plane, e));
    u128* rows = P.tables[G.table] + G.table_off + e * G.table_stride;

#pragma unroll 1
    for (uint32_t v = 0; v < G.q_in; ++v) {
        /* K_v = base_in + v * R_in: compress and hash it, then step to K_{v+1} */
        gc::Compressor ck(mi);
        uint32_t colour = 0;
        for (int d = (int)mi.nc - 1; d >= 0; --d) {
            const uint32_t dg = key[d * GB_THREADS];
            ck.push(dg);
            colour = dg;
            uint32_t nx = dg + __ldg(r_in + d);
            nx = nx >= mi.q ? nx - mi.q : nx;
            key[d * GB_THREADS] = (uint16_t)nx;
        }
        const u128 h = hash(aes, ck.value());
        if (G.mode == GB_MODE_DEC) { /* decoding information: H(K_v) at [v][element] */
            rows[(uint64_t)v * P.n + e] = h; /* table_stride is 0 for these */
            continue;
        }
        const uint32_t x = gb_function(G, v, r, G.q_out);
        if (G.mode == GB_MODE_MINI) {
            /* ProjectionGateMini::garble (projection_gate_mini.h:28-44): 16-bit slots of one row */
            reinterpret_cast<int16_t*>(rows)[colour] = (int16_t)(x + (uint32_t)(uint16_t)h.lo);
            continue;
        }
        gc::Compressor cp(mo);
        for (int d = (int)mo.nc - 1; d >= 0; --d) cp.push(gc::mod_q(base[d * GB_THREADS] + x * __ldg(off + d), mo));
        rows[colour] = gc::add128(cp.value(), h);
    }
}

cudaError_t launch_garble_gates(const GarbleParams& p, cudaStream_t s) {
    if (p.n == 0 || p.n_desc == 0) return cudaSuccess;
    static bool opted_in[MAX_DEVICES];
    const int dev = initialised_device();
    if (dev >= 0 && dev < MAX_DEVICES && !opted_in[dev]) {
        cudaError_t e = cudaFuncSetAttribute(k_garble_gates, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 128 * GB_THREADS * 2);
        if (e != cudaSuccess) return e;
        opted_in[dev] = true;
    }
    dim3 grid((unsigned)((p.n + GB_THREADS - 1) / GB_THREADS), p.n_desc);
    const size_t smem = 2 * (size_t)p.nc_max * GB_THREADS * sizeof(uint16_t);
    k_garble_gates<<<grid, GB_THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace dashb
