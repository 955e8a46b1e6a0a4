/*
 * kernels.h -- host-side launch interface of the sm_100a kernels
 * (kernels.cu) used by the C-ABI layer (dashb_api.cu).
 *
 * Device layout of a label tensor ("residue-major digit planes"): one plane
 * per CRT modulus j, u8 digits packed four to a 32-bit word,
 *     plane_j[element e][d/4][batch b][d%4]
 *         at  ((e*nc4_j + d/4)*Bp + b)*4 + d%4,   nc4_j = ceil(nc_j / 4),
 * Bp = batch rounded up to 16 (digits >= nc_j of the last word are zero).
 * A gadget thread (one element, one inference) moves 4 digits per 32-bit
 * access and a warp of 32 inferences 128 contiguous bytes; every element row
 * is 64-byte aligned and the whole plane is the K x N (N = 4*nc4_j*Bp)
 * row-major operand of the linear kernels as it stands.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "gc_common.cuh"

namespace dashb {

constexpr int MAXK = 12;   /* CRT moduli */
constexpr int MAXM = 12;   /* MRS digits */
constexpr int MAXF = 4;    /* rescale factors */

struct LinearParams {
    const uint8_t* in[MAXK];
    uint8_t* out[MAXK];
    const uint32_t* w4[MAXK];  /* [ceil(M/8)][Kp/4][8] packed weights mod q_j */
    const uint8_t* cst[MAXK];  /* [M][4*nc4_j] bias + zero-weight constant label, zero padded */
    gc::ModInfo mod[MAXK];
    const int32_t* koff;       /* [Kp] input element offset of tap k */
    uint32_t M, K, Kp;
    uint32_t v0, v1;           /* output rows [v0, v1) this launch computes (channel sharding) */
    uint32_t p0, p1;           /* output pixels [p0, p1) this launch computes (pixel sharding of a convolution) */
    uint32_t npix, ow;         /* output pixels per filter, output width */
    uint32_t step_x, step_y;   /* element offset per output x / y step */
    uint32_t Bp;
    int k;
    int exact_mod;             /* set by launch_linear: sums stay within the reciprocal's exact range */
    uint32_t y0;               /* set by launch_linear: first (pixel group, row group) index of this launch */
};

/* tcgen05 dense path (dense_tc.cu).  A tensor map is CUtensorMap's 128 opaque
 * bytes; kept opaque here so that this header does not need <cuda.h>. */
struct alignas(64) TensorMapBlob {
    uint64_t v[16];
};
struct DenseTcMaps {
    TensorMapBlob a[MAXK]; /* Wq_j  u8 [M][Kp16]            (K-major A)  */
    TensorMapBlob b[MAXK]; /* plane_j u8 [K][4*nc4_j*Bp]    (N-major B)  */
};
struct DenseTcParams {
    uint8_t* out[MAXK];
    const uint8_t* cst[MAXK];
    gc::ModInfo mod[MAXK];
    uint32_t M, K, Bp;
    uint32_t v0, v1;
    uint32_t nkb;          /* 128-column K blocks of the weight matrix */
    uint32_t b_box_bytes;  /* bytes one B box brings in (dense: 128*128) */
    /* convolution: per output pixel one GEMM; box {128, f_w, f_h, cb} at {n, x*sx, y*sy, kb*cb} */
    uint32_t conv, npix, ow, sx, sy, cb;
    uint32_t p0, p1;       /* output pixels [p0, p1) this launch computes (pixel sharding) */
    uint32_t npad;         /* != 0: filters on the N dimension (k_conv_tc_n), N = npad */
    /* strip != 0: k_conv_tc_strip -- a CTA computes `strip` horizontally adjacent output pixels from
     * strip + f_w - 1 shared input columns (stride 1 in x) */
    uint32_t strip;
    uint32_t st_rows;      /* K rows one input column contributes per channel block: cblk * f_h, a multiple of 32, <= 128 */
    uint32_t st_cblk;      /* channels per block */
    uint32_t st_ncb;       /* channel blocks */
    uint32_t st_fw;        /* filter width */
    uint32_t st_wboxes;    /* 128-row weight boxes per channel block: ceil(f_w * npad / 128) */
    int k;
};

/* element `le` (0 .. n-1) of a launch's slice: contiguous from e0, or -- a
 * convolution sharded by output pixels, whose elements are (filter, pixel) --
 * runs of `seg` pixels, one run per filter, `stride` = pixels per filter */
#if defined(__CUDACC__)
__host__ __device__
#endif
inline uint64_t slice_elem(uint64_t e0, uint64_t seg, uint64_t stride, uint64_t le) {
    return seg ? e0 + (le / seg) * stride + le % seg : e0 + le;
}

struct SignParams {
    const uint8_t* in[MAXK];
    uint8_t* out[MAXK];
    gc::ModInfo crt[MAXK];
    gc::ModInfo mrs[MAXM];
    gc::ModInfo cast[MAXM];  /* modulus (k+1)*m_i, i >= 1 */
    gc::ModInfo outm[MAXK];  /* sign layer: out moduli */
    gc::FracInfo castf[MAXM]; /* digit extraction plans of cast[] (gc_common.cuh) */
    gc::FracInfo crtf[MAXK];  /* ... of crt[] */
    const gc::u128 *approx, *cast1, *cast2, *sign, *garbler, *evaluator;
    gc::u128 carry0;         /* compressed zero label of m_{n_mrs-1} */
    uint64_t n;   /* elements this launch processes ... */
    uint64_t e0;  /* ... starting at this element (channel sharding) */
    uint64_t seg, stride; /* seg != 0: the n elements are runs of seg elements, run s starting at e0 + s * stride (slice_elem) */
    uint32_t prefix[MAXK];
    uint32_t S, Mp;
    uint32_t batch, Bp;
    int k, n_mrs, n_out;
    int lanes;               /* sub-batch lanes evaluating concurrently (picks latency vs throughput kernel) */
    uint32_t tab_mul, tab_add; /* element e uses the tables of gadget instance e * tab_mul + tab_add (1, 0 except max pooling) */
    /* rep > 1: every table row exists rep times, row r of inference b at r * rep + (b0 + b) % rep -- the layout a
     * garbling per inference gives a batch (one set of tables per inference, rows interleaved by inference) */
    uint32_t rep, b0;
};

/* k_pool_step (max pooling, max(x, y) = x + relu(y - x)): per output element
 * and kernel position one ReLU gadget; the label arithmetic around it */
struct PoolParams {
    const uint8_t* in[MAXK];  /* the layer's input planes */
    uint8_t* cur[MAXK];       /* running maximum = the layer's output planes */
    uint8_t* diff[MAXK];      /* y - x, input of the ReLU gadget */
    const uint8_t* relu[MAXK];/* its output */
    gc::ModInfo crt[MAXK];
    uint32_t in_w, in_h, ch, ow, oh, s_w, s_h;
    uint32_t l, m;            /* kernel position of this step */
    uint64_t n_out;
    uint32_t Bp;
    int k;
    int mode;                 /* 0: cur = in[first]; diff = in[pos] - cur   1: cur += relu; diff = in[pos] - cur   2: cur += relu */
};

/* k_ingest_direct: labels read straight from the caller's (pinned, mapped)
 * host buffers into the digit planes, all moduli in one launch */
struct IngestParams {
    const int16_t* src[MAXK]; /* device-visible address of the lane's first inference, int16 [batch][n][nc_j] */
    const int16_t* end[MAXK]; /* one past the last digit the lane may read */
    uint8_t* plane[MAXK];
    uint32_t nc[MAXK];
    uint64_t n;
    uint32_t batch, Bp;
    int k;
};

/* k_egress_all: the output planes of every modulus back to int16 [batch][n][nc_j] in one launch */
struct EgressParams {
    const uint8_t* plane[MAXK];
    int16_t* dst[MAXK]; /* the lane's first inference; device memory or mapped pinned host memory */
    uint32_t nc[MAXK];
    uint64_t n;
    uint32_t batch, Bp;
    int k;
};

/* k_shard_push: one rank's slice of an activation tensor stored into the
 * gather buffer of every rank (its own included) over NVLink peer memory */
constexpr int MAXR = 16; /* ranks of an output-channel-sharded evaluation */
struct ShardPushParams {
    const uint8_t* src[MAXK]; /* plane j of the tensor the layer group produced */
    uint64_t dst_off[MAXK];   /* offset of plane j inside a gather buffer */
    uint64_t row[MAXK];       /* bytes per element of plane j */
    uint64_t e0, e1;          /* this rank's elements: [e0, e1), or with seg != 0 ... */
    uint64_t seg, stride, nseg; /* ... nseg runs of seg elements, run s starting at e0 + s * stride */
    uint8_t* peer[MAXR];      /* every rank's gather buffer (this rank's own at [rank]) */
    int k, world;
};

/* k_shard_push_comp / k_shard_unpack: the same exchange with 16-byte compressed labels (shard_kernels.cuh) */
struct ShardCompParams {
    const uint8_t* plane[MAXK]; /* the lane's current tensor: read by the push, rewritten by the unpack */
    gc::ModInfo mod[MAXK];
    uint8_t* peer[MAXR];  /* every rank's gather buffer: u128 [modulus][element][inference] */
    const uint8_t* local; /* this rank's own gather buffer */
    uint64_t n;           /* elements of the tensor */
    uint64_t e0, seg, stride, count; /* this rank's slice: `count` elements, see slice_elem */
    uint32_t Bp;
    int k, world;
};

struct ProjParams {
    const uint8_t* in[MAXK];
    uint8_t* out[MAXK];
    gc::ModInfo min[MAXK], mout[MAXK];
    const gc::u128* ciphers[MAXK];
    uint64_t n;   /* elements this launch processes ... */
    uint64_t e0;  /* ... starting at this element (channel sharding) */
    uint64_t seg, stride; /* see SignParams */
    uint32_t batch, Bp;
    int k;
};

struct RescaleParams {
    const uint8_t* in[MAXK];
    uint8_t* out[MAXK];
    gc::ModInfo crt[MAXK];
    gc::FracInfo crtf[MAXK];    /* digit extraction plans of crt[] (gc_common.cuh) */
    const uint8_t* up[MAXK];    /* upshift label digits (u8[4*nc4_j], zero padded) */
    const uint8_t* down[MAXK];  /* downshift label digits */
    const uint8_t* zero[MAXK];  /* zero label digits */
    uint32_t zstride[MAXK];     /* 0: zero[j] is one label for every element; else element e's label starts at zero[j] + e * zstride[j]
                                 * (standalone base extension: the labels the layer starts the extra residues from) */
    uint32_t loff[MAXK];        /* (unused by the kernel) running offset of residue j's digits */
    const gc::u128* tm;
    const gc::u128* be;
    uint64_t tm_off[MAXF][MAXK]; /* row offset of block (factor i, modulus j) */
    uint16_t inv_f[MAXF][MAXK];  /* factor_i^-1 mod p_j */
    uint32_t factor[MAXF];
    int fidx[MAXF];
    int perm[MAXK];              /* swapped order -> CRT index */
    uint64_t be_off[MAXK][MAXK]; /* row offset of BE block (i, t) in swapped order */
    uint16_t inv_partial[MAXK][MAXK];
    uint16_t neg_invv[MAXF];     /* (-invv) mod extra modulus */
    int extra_idx[MAXF];         /* CRT index of extra modulus x */
    int extra_pos[MAXF];         /* its position in the swapped order */
    uint64_t n;   /* elements this launch processes ... */
    uint64_t e0;  /* ... starting at this element (channel sharding) */
    uint64_t seg, stride; /* see SignParams */
    uint32_t batch, Bp;
    int k, nf, non_ext;
    int legacy;                  /* 1: stop after step 2.3 (sign gadget and downshift follow) */
    /* lazy-reduction constants (see k_rescale): a chain of updates
     * w <- (w - p_i) * inv_i is w0*a + sum_i p_i*b_i (mod q) */
    uint16_t a1[MAXK];           /* step 2 on residue j: product of the factors' inverses */
    uint16_t b1[MAXF][MAXK];     /* coefficient of factor i's projection into residue j */
    uint16_t a1f[MAXF];          /* same for factor i's own residue (earlier factors only) */
    uint16_t b1f[MAXF][MAXF];
    uint16_t a2[MAXK];           /* base extension, target t (swapped order) */
    uint16_t b2[MAXK][MAXK];     /* coefficient of source i's projection into target t */
    int wide;                    /* 1: some (k+1)(q-1)^2 >= 2^16, use 32-bit accumulators */
};

struct AddLabelParams {
    uint8_t* plane[MAXK];
    const uint8_t* label[MAXK]; /* u8[4*nc4_j], zero padded */
    gc::ModInfo crt[MAXK];
    uint64_t n;   /* elements this launch processes ... */
    uint64_t e0;  /* ... starting at this element (channel sharding) */
    uint32_t Bp;
    int k;
    int subtract;
};

struct GarbleInParams {
    const int64_t* x;           /* [batch][n] plaintext inputs of this sub-batch */
    uint8_t* out[MAXK];
    const uint8_t* base[MAXK];  /* base labels u8 [n][4*nc4_j], zero padded */
    const uint8_t* off[MAXK];   /* offset label R_{p_j} u8 [4*nc4_j] */
    gc::ModInfo crt[MAXK];
    uint64_t n;
    uint32_t batch, Bp;
    int k;
};

struct DecodeParams {
    const uint8_t* in[MAXK];
    gc::ModInfo crt[MAXK];
    const gc::u128* dec[MAXK]; /* [p_j][n] */
    int64_t inv[MAXK];         /* mul_inv(prod / p_j, p_j) mod prod */
    int64_t pdiv[MAXK];        /* prod / p_j */
    int64_t prod, crt_modulus;
    int64_t* values;           /* [batch][n] */
    uint8_t* ok;               /* [batch] */
    uint64_t n;
    uint32_t batch, Bp;
    int k;
};

/* ---- GPU garbler (garble_kernels.cuh, garbler.cu) ---- */
constexpr uint32_t GB_NONE = 0xffffffffu;
enum { GB_TERM_PLANE = 1, GB_TERM_CONST = 2, GB_TERM_PRF = 3 };
enum { GB_F_ID = 0, GB_F_LOOKUP, GB_F_DIV, GB_F_SIGN, GB_F_MULR, GB_F_NEGSUM, GB_F_ADDR };
enum { GB_MODE_PROJ = 0, GB_MODE_MINI = 1, GB_MODE_DEC = 2 };
constexpr int GB_MAXTERMS = 14;

/* one summand of a label: coef * (label of input plane `arg` | constant label
 * at byte offset `arg` of GarbleParams::labels | PRF label with tag `arg`) */
struct GateTerm {
    uint32_t kind, coef;
    uint64_t arg;
};

/* one projection gate (or, with q_in == 0, just an output label) per element */
struct GateDesc {
    uint32_t q_in, q_out;
    uint32_t n_in, n_out;
    GateTerm in[GB_MAXTERMS];   /* the gate's input base label, modulus q_in */
    GateTerm out[GB_MAXTERMS];  /* its output base label, modulus q_out */
    uint32_t off_plane;         /* payload offset label: GB_NONE = R_{q_out}, else that input plane's label */
    uint32_t r_plane;           /* r = colour digit of that input plane's label (half gates), GB_NONE: 0 */
    uint32_t f;                 /* GB_F_* */
    int32_t fa, fb, fc;
    uint64_t flut;              /* GB_F_LOOKUP: device pointer, value of v at flut[v * flut_stride] */
    uint32_t flut_stride;
    uint32_t mode;              /* GB_MODE_* */
    uint32_t table;             /* index into GarbleParams::tables */
    int32_t out_plane;          /* >= 0: store the output label to that output plane */
    uint64_t table_off, table_stride; /* rows at table_off + element * table_stride (+ colour) */
};

struct GarbleParams {
    const GateDesc* desc;
    uint32_t n_desc;
    uint32_t nc_max;            /* longest label (digits) any descriptor touches */
    uint64_t n;                 /* elements */
    const uint8_t* in[MAXK];    /* input base labels, u8 [element][4*nc4] */
    uint32_t in_nc[MAXK];
    uint8_t* out[MAXK];         /* output base labels, same layout */
    gc::u128* tables[16];
    const gc::ModInfo* mods;    /* indexed by modulus */
    const uint8_t* labels;      /* constant labels: u8 digits, addressed by byte offset */
    const uint32_t* r_off;      /* byte offset of the offset label R_q, indexed by modulus */
    uint64_t layer_tag;
};

cudaError_t garbler_set_prf_key(const uint32_t rk[44]);
cudaError_t launch_garble_gates(const GarbleParams& p, cudaStream_t s);

/* small services of dashb_api.cu for entry points that live next to their
 * kernels (dashb_projection_static) */
gc::ModInfo modinfo_for(uint32_t q);
int report_error(int code, const char* fmt, ...);
cudaStream_t library_stream();
int initialised_device();

/* uploads T-tables / S-box / round keys of key 00..0f to the device */
cudaError_t upload_aes_tables();
/* host copies of the tables for dashb_hostcheck_* */
void host_aes_tables(const uint32_t** te0, const uint32_t** te1, const uint32_t** te2, const uint32_t** te3,
                     const uint8_t** sbox, const uint32_t** rk);

cudaError_t launch_ingest(const int16_t* src, uint8_t* plane, uint64_t n, uint32_t nc, uint32_t batch, uint32_t Bp,
                          cudaStream_t s);
cudaError_t launch_ingest_direct(const IngestParams& p, cudaStream_t s);
cudaError_t launch_egress(const uint8_t* plane, int16_t* dst, uint64_t n, uint32_t nc, uint32_t batch, uint32_t Bp,
                          cudaStream_t s);
cudaError_t launch_ingest_compressed(const void* src, uint8_t* plane, uint64_t n, const gc::ModInfo& m, uint32_t batch, uint32_t Bp,
                                     cudaStream_t s);
cudaError_t launch_egress_compressed(const uint8_t* plane, void* dst, uint64_t n, const gc::ModInfo& m, uint32_t batch, uint32_t Bp,
                                     cudaStream_t s);
cudaError_t launch_linear(const LinearParams& p, cudaStream_t s);
cudaError_t dense_tc_make_map(void* map_out, const void* ptr, uint64_t inner, uint64_t outer);
cudaError_t conv_tc_make_map(void* map_out, const void* ptr, uint64_t row_bytes, uint64_t in_w, uint64_t in_h,
                             uint64_t channels, uint32_t f_w, uint32_t f_h, uint32_t cb);
cudaError_t launch_dense_tc(const DenseTcMaps& maps, const DenseTcParams& p, cudaStream_t s);
cudaError_t launch_sign(const SignParams& p, bool relu, cudaStream_t s);
cudaError_t launch_egress_all(const EgressParams& p, cudaStream_t s);
cudaError_t launch_projection(const ProjParams& p, cudaStream_t s);
cudaError_t launch_shard_push(const ShardPushParams& p, cudaStream_t s);
cudaError_t launch_shard_push_comp(const ShardCompParams& p, cudaStream_t s);
cudaError_t launch_shard_unpack(const ShardCompParams& p, cudaStream_t s);
cudaError_t launch_shard_signal(uint32_t* const* peer_flags, int world, int rank, uint32_t epoch, cudaStream_t s);
cudaError_t launch_shard_wait(uint32_t* flags, int world, uint32_t epoch, cudaStream_t s);
cudaError_t launch_pool_step(const PoolParams& p, cudaStream_t s);
cudaError_t launch_replicate_rows(const void* in, void* out, uint64_t rows, uint32_t rep, cudaStream_t s);
cudaError_t launch_rescale(const RescaleParams& p, cudaStream_t s);
cudaError_t launch_add_label(const AddLabelParams& p, cudaStream_t s);
cudaError_t launch_decode(const DecodeParams& p, cudaStream_t s);
cudaError_t launch_garble_inputs(const GarbleInParams& p, cudaStream_t s);

}  // namespace dashb
