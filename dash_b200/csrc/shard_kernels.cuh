/*
 * shard_kernels.cuh -- the compressed form of the peer-memory exchange of an
 * output-sharded evaluation, included at the end of
 * kernels.cu so that it shares load_compress / store_digits.
 *
 * A label crosses NVLink as the 16 bytes of LabelTensor::compress
 * (garbling/label_tensor.h:715-727) instead of its 4*nc4 digit bytes (k = 7
 * scaling base: 112 instead of 332 bytes per element and inference).
 * k_shard_push_comp compresses this rank's slice of the lane's tensor and
 * stores it into every rank's gather buffer ([modulus][element][inference]
 * u128); after the epoch wait, k_shard_unpack expands every element of the
 * tensor back into the lane's digit planes.  Compression is a bijection on
 * valid labels, so the planes come out byte for byte as they were pushed.
 * Plain loads and stores only.
 */
namespace dashb {

__global__ void __launch_bounds__(128) k_shard_push_comp(const __grid_constant__ ShardCompParams P) {
    const int j = blockIdx.y;
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.count * P.Bp) return;
    const uint32_t b = (uint32_t)(idx % P.Bp);
    const uint64_t e = slice_elem(P.e0, P.seg, P.stride, idx / P.Bp);
    const ModInfo m = pin(P.mod[j]);
    uint32_t colour;
    const u128 v = load_compress(P.plane[j], m, e, b, P.Bp, colour);
    const uint4 w = make_uint4((uint32_t)v.lo, (uint32_t)(v.lo >> 32), (uint32_t)v.hi, (uint32_t)(v.hi >> 32));
    const uint64_t at = (((uint64_t)j * P.n + e) * P.Bp + b) * 16;
    for (int r = 0; r < P.world; ++r) *reinterpret_cast<uint4*>(P.peer[r] + at) = w;
    __threadfence_system();
}

/* every element of the tensor, this rank's own slice included (it costs 1/world of the work and keeps the
 * kernel free of slice arithmetic) */
__global__ void __launch_bounds__(128) k_shard_unpack(const __grid_constant__ ShardCompParams P) {
    const int j = blockIdx.y;
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.n * P.Bp) return;
    const uint32_t b = (uint32_t)(idx % P.Bp);
    const uint64_t e = idx / P.Bp;
    const ModInfo m = pin(P.mod[j]);
    const uint4 w = *reinterpret_cast<const uint4*>(P.local + (((uint64_t)j * P.n + e) * P.Bp + b) * 16);
    u128 v;
    v.lo = (uint64_t)w.x | ((uint64_t)w.y << 32);
    v.hi = (uint64_t)w.z | ((uint64_t)w.w << 32);
    store_digits(const_cast<uint8_t*>(P.plane[j]), m, e, b, P.Bp, v);
}

cudaError_t launch_shard_push_comp(const ShardCompParams& p, cudaStream_t s) {
    const uint64_t total = p.count * p.Bp;
    if (total == 0) return cudaSuccess;
    dim3 grid((unsigned)((total + 127) / 128), p.k);
    k_shard_push_comp<<<grid, 128, 0, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_shard_unpack(const ShardCompParams& p, cudaStream_t s) {
    const uint64_t total = p.n * p.Bp;
    if (total == 0) return cudaSuccess;
    dim3 grid((unsigned)((total + 127) / 128), p.k);
    k_shard_unpack<<<grid, 128, 0, s>>>(p);
    return cudaGetLastError();
}

}  // namespace dashb
