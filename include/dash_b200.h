/*
 * dash_b200.h -- C ABI of the B200-native evaluator for Dash's online
 * garbled-inference path.
 *
 * Everything here is `extern "C"`, plain pointers and sizes.  The library
 * owns all device memory; the caller owns every host buffer it passes in.
 * Every function returns DASHB_OK (0) or a negative DASHB_E* code and never
 * throws; dashb_last_error() returns a thread-local description of the last
 * failure.  There is no CPU fallback: without a CUDA device dashb_init()
 * fails with DASHB_ENODEV and every later call fails too.
 *
 * Citations are file:line in UzL-ITS/dash (dash/include/... unless stated).
 * The reference's own out-of-class operator boundary is the SGX ocall set
 * (sgx/Enclave/Enclave.edl:52-134); the entry points below are that set
 * generalised with a batch dimension and completed with the operators the
 * reference never moved to the GPU (general rescale, base extension).
 *
 * Label layout at this boundary (identical to LabelTensor::get_components(),
 * garbling/label_tensor.h:238-246): for CRT modulus j a host buffer of
 * int16 `[batch][label][nc(p_j)]`, digit 0 = colour, digits in [0, p_j).
 * nc(q) = floor(128 / log2 q) (label_tensor.h:1353-1357).
 * The batch dimension and security.  A batch evaluates ONE garbled circuit
 * (one set of tables, one set of wire labels) on several inputs.  The
 * reference has no batch: one garbling serves one inference, and that is
 * what the scheme's security needs -- an evaluator who sees two labels of the
 * same wire learns (x - x') * R_q and from it the global offset labels.
 * Batching several inputs through one circuit is therefore only sound for
 * testing and benchmarking, or when all inputs of the batch come from one
 * party that accepts them being linked; independent private inferences each
 * need their own garbling (dashb_garble_to_circuit makes one in milliseconds)
 * and their own circuit object.  Throughput numbers quoted for a shared-table
 * batch are an upper bound for that deployment: the kernels do the same work
 * per inference, but the tables then stream from HBM instead of staying in L2.
 *
 * Ciphertext tables are arrays of 16-byte little-endian unsigned integers
 * (`__uint128_t` in the reference) in exactly the order the reference's
 * garbler fills them; the per-operator comments give that order.
 */
#ifndef DASH_B200_H
#define DASH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DASHB_OK 0
#define DASHB_EINVAL (-1)   /* bad argument / unsupported configuration      */
#define DASHB_ENODEV (-2)   /* no CUDA device / library used before init     */
#define DASHB_ECUDA (-3)    /* a CUDA runtime call or kernel failed          */
#define DASHB_ESTATE (-4)   /* call order violated (e.g. add after finalize) */
#define DASHB_EIO (-5)      /* material file could not be read / written     */
#define DASHB_ENOMEM (-6)   /* device or host allocation failed              */

/* One ciphertext row: the reference's __uint128_t, little endian. */
typedef struct dashb_u128 {
    uint64_t lo, hi;
} dashb_u128;

typedef struct dashb_circuit dashb_circuit; /* opaque */

/* ---- process / device ------------------------------------------------- */

/* Selects `device`, uploads the AES tables and the round keys of the fixed
 * key 00..0f.  Replaces init_cuda() (misc/cuda_util.h:19-23) and the
 * CUDAAESEngine constructor (crypto/cuda_aes_engine.h:533-545).  One process
 * drives one GPU; call once per process before anything else. */
int dashb_init(int device);
const char* dashb_last_error(void);
/* nc(q) exactly as LabelTensor::get_nr_comps (label_tensor.h:1353-1357). */
int dashb_nr_comps(int modulus);
/* The CUDA stream all work of this library is enqueued on (a cudaStream_t
 * cast to void*), so a caller can bracket calls with its own events. */
void* dashb_stream(void);

/* ---- circuit construction (replaces GarbledCircuitInterface::cuda_move,
 *      garbled_circuit_interface.h:477-549, and each layer's cuda_move) --- */

/* crt_base[k], mrs_base[n_mrs] (n_mrs may be 0 for circuits without
 * sign/ReLU/legacy-rescale layers); zero_labels[q-2] = int16[nc(q)] for
 * q = 2..max_modulus (get_zero_labels(), garbled_circuit_interface.h:768);
 * upshift_labels[j] = int16[nc(p_j)] (get_upshift_label, :786) or NULL when
 * the circuit has no rescale layer; input_size = number of circuit inputs. */
int dashb_circuit_create(dashb_circuit** out, const int32_t* crt_base, int k,
                         const int32_t* mrs_base, int n_mrs, int max_modulus,
                         const int16_t* const* zero_labels,
                         const int16_t* const* upshift_labels,
                         uint64_t input_size);

/* GarbledDense (layer/garbled_dense.h:120-144,189-223; kernel
 * MatMulModAddZero[TF], misc/cuda_util.h:95-138; ocall_cuda_eval_dense,
 * Enclave.edl).  weights = get_qe_weights(): int64 [M][K] reduced to [0,P);
 * bias_labels[j] = int16 [M][nc(p_j)]; channel_tf as Dense::get_channel_tf
 * (0 = plain, >0 = TF channel-last input permutation). */
int dashb_add_dense(dashb_circuit* c, uint64_t M, uint64_t K, int channel_tf,
                    const int64_t* weights,
                    const int16_t* const* bias_labels);

/* GarbledConv2d (layer/garbled_conv2d.h:113-143; CPU semantics of
 * LabelTensor::conv2d, label_tensor.h:920-1002, including its weight index
 * v*fs*C + w*fs + i*filter_height + j).  weights int64
 * [filters*channels*f_w*f_h] reduced to [0,P); bias_labels[j] = int16
 * [filters][nc(p_j)]. */
int dashb_add_conv2d(dashb_circuit* c, uint64_t in_w, uint64_t in_h,
                     uint64_t channels, uint64_t filters, uint64_t f_w,
                     uint64_t f_h, uint64_t s_w, uint64_t s_h,
                     const int64_t* weights,
                     const int16_t* const* bias_labels);

/* GarbledRelu (layer/garbled_relu.h:181-223,258-294) = SignGadget with out
 * modulus {2} and bounds (0,1) + one MixedModHalfGate per CRT modulus.
 * With S = sum(crt_base), M' = sum(mrs_base[1:]):
 *   approx    [n][n_mrs*S]     (sign_gadget.h:295-317; per element, CRT
 *                               modulus j at n_mrs*prefix_j, MRS digit i at
 *                               + i*p_j, row = colour)
 *   cast1     [n][(k+1)*M']    (sign_gadget.h:489-529; per element, digits
 *                               n_mrs-1..1, per digit k+1 tables of m_i rows)
 *   cast2     [n][(k+1)*M']    (sign_gadget.h:443-445,534-552; per element,
 *                               digits n_mrs-1..1, one table of (k+1)*m_i)
 *   sign      [n][1][m_0]      (sign_gadget.h:557-576)
 *   garbler   [n][S]           (garbled_relu.h:158)
 *   evaluator [n][k][3]        (garbled_relu.h:160; rows 0,1 = projection
 *                               2->p_j, row 2 = packed int16 mini gate,
 *                               gates/projection_gate_mini.h:40-41) */
int dashb_add_relu(dashb_circuit* c, uint64_t n, const dashb_u128* approx,
                   const dashb_u128* cast1, const dashb_u128* cast2,
                   const dashb_u128* sign, const dashb_u128* garbler,
                   const dashb_u128* evaluator);

/* GarbledMaxPool2d (layer/garbled_maxpool2d.h:143-224; CPU-only in the
 * reference, :226-244) = per output (i, j, channel) a chain of MaxGadgets
 * (gadgets/max_gadget.h:116-160: max(x, y) = x + relu(y - x)) over the kernel
 * positions (l, m), l outer, (0, 0) included, each with a ReluGadget of its
 * own (gadgets/relu_gadget.h).  The six tables are those of dashb_add_relu for
 *   n = outputs * k_w * k_h  gadget instances, instance of (output o, l, m)
 *   = o * k_w * k_h + l * k_h + m,  o = (i * oh + j) * channels + channel
 *   (the output label's index; this layer counts the last index fastest on
 *   both sides -- input (x, y, z) at (x * in_h + y) * channels + z,
 *   label_tensor.h:1574-1579 and 1290-1301 -- unlike conv2d):
 * the concatenation of every instance's one-element SignGadget / half-gate
 * arrays in that order.  Geometry is passed explicitly; the reference layer
 * reads its input dims from the circuit's input labels (:66-68), so it is
 * only defined as the first layer and for shapes where its window count
 * ((W - k_w) / s_w + 1) equals its output tensor's W / k_w (max_pool2d.h:
 * 26-29) -- other shapes are rejected. */
int dashb_add_maxpool2d(dashb_circuit* c, uint64_t in_w, uint64_t in_h,
                        uint64_t channels, uint64_t k_w, uint64_t k_h,
                        uint64_t s_w, uint64_t s_h, const dashb_u128* approx,
                        const dashb_u128* cast1, const dashb_u128* cast2,
                        const dashb_u128* sign, const dashb_u128* garbler,
                        const dashb_u128* evaluator);

/* GarbledSign (layer/garbled_sign.h:16-81): SignGadget, bounds (-1,1), out
 * moduli = crt_base; sign table is [n][k][m_0]. */
int dashb_add_sign(dashb_circuit* c, uint64_t n, const dashb_u128* approx,
                   const dashb_u128* cast1, const dashb_u128* cast2,
                   const dashb_u128* sign);

/* GarbledRescale with general factors + BaseExtensionGadget
 * (layer/garbled_rescale.h:75-103, gadgets/rescale_gadget.h:244-362,
 * gadgets/base_extension_gadget.h:197-295; CPU-only in the reference).
 * factors[nf] are CRT moduli; downshift_labels[j] = int16[nc(p_j)]
 * (get_downshift_label, garbled_circuit_interface.h:803-815);
 *   trans_mod  sequential [factor i][modulus j not yet ignored][element]
 *              [s_i rows]                      (rescale_gadget.h:173-175)
 *   be_ciphers sequential [i][j][element][m_i rows] over the swapped base
 *              (base_extension_gadget.h:120-152). */
int dashb_add_rescale(dashb_circuit* c, uint64_t n, const int32_t* factors,
                      int nf, const int16_t* const* downshift_labels,
                      const dashb_u128* trans_mod,
                      const dashb_u128* be_ciphers);

/* GarbledBaseExtension (layer/garbled_base_extension.h:67-97; CPU-only and
 * "just for testing purposes" in the reference, circuit/layer/
 * base_extension.h:10-14): the residues of the `extra_moduli` (CRT moduli) are
 * rebuilt from the other residues by BaseExtensionGadget::cpu_evaluate
 * (gadgets/base_extension_gadget.h:197-295); the other residues pass through
 * unchanged.  The reference first REPLACES each extra residue's incoming label
 * by the circuit's input base label of the same index (:74-80):
 * start_labels[x] = int16 [n][nc(extra_moduli[x])] are those labels
 * (get_base_label().at(index of the modulus)).  be_ciphers as in
 * dashb_add_rescale. */
int dashb_add_base_extension(dashb_circuit* c, uint64_t n, const int32_t* extra_moduli,
                             int n_extra, const int16_t* const* start_labels,
                             const dashb_u128* be_ciphers);

/* Legacy GarbledRescale{l}: l iterations of factor 2, each followed by a
 * SignGadget with bounds (1,0), out modulus {2} that regenerates residue 0
 * (layer/garbled_rescale.h:27-35,93-100; rescale_gadget.h:208-224,328-342).
 * Requires crt_base[0] == 2.  Arrays are per iteration it < l:
 * trans_mod[it] [(k-1)][n][2], then the four SignGadget tables as in
 * dashb_add_relu (sign [n][1][m_0]). */
int dashb_add_rescale_legacy(dashb_circuit* c, uint64_t n, int l,
                             const int16_t* const* downshift_labels,
                             const dashb_u128* const* trans_mod,
                             const dashb_u128* const* approx,
                             const dashb_u128* const* cast1,
                             const dashb_u128* const* cast2,
                             const dashb_u128* const* sign);

/* GarbledProjection (layer/garbled_projection.h:118-134,168-194; kernel
 * eval_proj, misc/cuda_util.h:297-317): ciphers[j] = [n][in_moduli[j]].
 * The output labels live in out_moduli[j], which become the moduli of the
 * following layer's input. */
int dashb_add_projection(dashb_circuit* c, uint64_t n,
                         const int32_t* in_moduli, const int32_t* out_moduli,
                         const dashb_u128* const* ciphers);

/* Projection gates without any allocation.  No counterpart in the
 * reference.
 *
 * For hosts that must not allocate while they evaluate: a server that gave
 * its whole device-memory budget to its own arena at start-up, the untrusted
 * half of Dash's SGX split where every cudaMalloc is an ocall round trip out
 * of the enclave (sgx/Enclave/Enclave.edl, ocall_cudaMalloc), or a latency-
 * critical loop in which cudaMalloc/cudaFree would stall the whole device.
 * Such a caller keeps its labels and tables in device memory it already owns
 * and passes plain device pointers; the digit planes the kernels need in
 * between live in one fixed array inside the library image
 * (DASHB_STATIC_SCRATCH_BYTES, resident on every GPU that loads the library
 * until the process exits).  The call itself performs no cudaMalloc, no
 * cudaFree and calls no allocating library routine.
 *
 * Evaluates n projection gates of one modulus pair for `batch` inferences:
 *   d_in      device int16 [batch][n][nc(q_in)]   (LabelTensor digit order)
 *   d_ciphers device 16-byte rows [n][q_in]
 *   d_out     device int16 [batch][n][nc(q_out)]
 * i.e. GarbledProjection / eval_proj semantics, identical to what
 * dashb_evaluate computes for a projection layer.  q_in, q_out in [2,256].
 * Work is enqueued on `stream` (a cudaStream_t; NULL = dashb_stream()) and
 * is asynchronous.  Fails with DASHB_EINVAL, touching nothing, when
 *   4*Bp*n*(ceil(nc(q_in)/4) + ceil(nc(q_out)/4)) > DASHB_STATIC_SCRATCH_BYTES
 * (Bp = batch rounded up to 16).
 *
 * The first call on a device looks up the array's device address (one-off
 * set-up); make that first call before you start any stream capture.
 * All calls share the one array and it keeps whatever the previous call left:
 * the part a call uses is cleared at its start, so results never depend on
 * earlier calls, but two calls that overlap in time on the GPU (different
 * streams, or several host threads) race on it -- serialise them.
 * Requires dashb_init() on the current device. */
#define DASHB_STATIC_SCRATCH_BYTES (8u << 20)
int dashb_projection_static(const int16_t* d_in, const dashb_u128* d_ciphers, int16_t* d_out, uint64_t n,
                            uint64_t batch, int q_in, int q_out, void* stream);

/* Decoding information (gen_decoding_information,
 * garbled_circuit_interface.h:386-415): dec[j] = [p_j][output_size] hashes.
 * Optional; needed only for dashb_decode_outputs. */
int dashb_set_decoding_info(dashb_circuit* c, const dashb_u128* const* dec);

/* Garbler-side input encoding for deployments where the evaluating process
 * also encodes the inputs (benchmarks, single-party tests): base_labels[j] =
 * int16 [input_size][nc(p_j)] (get_base_label(),
 * garbled_circuit_interface.h:771) and offset_labels[j] = int16 [nc(p_j)],
 * the global offset R_{p_j} (get_label_offset, :762-767).  Optional; needed
 * only for dashb_garble_inputs.  Call before dashb_finalize. */
int dashb_set_input_encoding(dashb_circuit* c, const int16_t* const* base_labels,
                             const int16_t* const* offset_labels);

/* The memory behaviour of one garbling per inference, for measurement.  A
 * batch normally shares one circuit's ciphertext tables (see the note on the
 * batch dimension above), which keeps them L2-resident.  With replicas = R
 * (before the first layer is added) every sign / ReLU table is stored R times,
 * row r of inference b at r * R + b % R -- the layout R independent garblings
 * would have for a batch (rows interleaved by inference), here filled with R
 * copies of the one garbling, so results are unchanged and checkable while
 * every inference streams its own rows from HBM.  Labels and linear-layer
 * constants stay shared; rescale / max-pooling / legacy-rescale tables are
 * not replicated (the latter two reject replicas).  bench.py reports the
 * throughput in this mode next to the shared-table number. */
int dashb_set_table_replicas(dashb_circuit* c, int replicas);

/* Uploads weights and tables, re-lays them for the kernels and sizes the
 * activation buffers for up to max_batch concurrent inferences. */
int dashb_finalize(dashb_circuit* c, uint64_t max_batch);

/* ---- online evaluation ------------------------------------------------ */

/* cuda_move_inputs (garbled_circuit_interface.h:551-565), batched:
 * in[j] = int16 [batch][input_size][nc(p_j)]. */
int dashb_move_inputs(dashb_circuit* c, const int16_t* const* in,
                      uint64_t batch);
/* garble_inputs (garbled_circuit_interface.h:325-343) on the device,
 * batched: label_j[e] = (x[e] mod p_j) * R_{p_j} + base_j[e] digit-wise mod
 * p_j, written straight into the input planes; values = host int64
 * [batch][input_size].  Replaces garble_inputs + cuda_move_inputs: only
 * 8 bytes per input value cross PCIe.  Needs dashb_set_input_encoding. */
int dashb_garble_inputs(dashb_circuit* c, const int64_t* values, uint64_t batch);

/* cuda_evaluate (garbled_circuit_interface.h:603-612): runs every layer on
 * the labels left by dashb_move_inputs; asynchronous on dashb_stream(). */
int dashb_evaluate_device(dashb_circuit* c);
/* cuda_move_outputs (garbled_circuit_interface.h:733-736), batched:
 * out[j] = int16 [batch][output_size][nc(q_j)], q_j = modulus j of the last
 * layer's output.  Synchronises. */
int dashb_move_outputs(dashb_circuit* c, int16_t* const* out);
/* The three calls above in sequence. */
int dashb_evaluate(dashb_circuit* c, const int16_t* const* in, uint64_t batch,
                   int16_t* const* out);
/*
 * dashb_evaluate for a caller that streams batches: returns as soon as the
 * step is enqueued; up to two steps are in flight (the lanes' streams run the
 * steps in order, the input planes are double-buffered), so that step n+1's
 * input labels cross PCIe while step n computes (the third call first waits
 * for the oldest step).  Output buffers in pinned, mapped host memory are
 * written by the egress kernel directly; pageable ones through a staging copy.  `in` and `out` must stay valid and
 * untouched until that step is done: after dashb_evaluate_wait(), or after
 * the second-next dashb_evaluate_async returns.  Steps complete in order.
 * The work is NOT ordered with dashb_stream(); dashb_move_outputs /
 * dashb_decode_outputs / dashb_layer_outputs refer to synchronous
 * evaluations only (any synchronous call first drains the asynchronous
 * ones).  No counterpart in the reference, whose GPU calls are synchronous
 * (garbled_circuit_interface.h:551-612,733-736).
 */
int dashb_evaluate_async(dashb_circuit* c, const int16_t* const* in, uint64_t batch, int16_t* const* out);
int dashb_evaluate_wait(dashb_circuit* c);
/*
 * The same three calls with labels in their compressed 128-bit form on the
 * host side (what LabelTensor::compress() leaves in get_compressed(),
 * garbling/label_tensor.h:715-727; SURVEY.md row f2: "compressed 128-bit
 * labels are the natural wire form"): per CRT modulus dashb_u128
 * [batch][label].  16 bytes per label cross PCIe instead of 2*nc (6.5x less
 * for the k = 9 base); the digits are extracted / recombined on the device.
 * No counterpart in the reference, whose GPU path always copies digit arrays.
 */
int dashb_move_inputs_compressed(dashb_circuit* c, const dashb_u128* const* in, uint64_t batch);
int dashb_move_outputs_compressed(dashb_circuit* c, dashb_u128* const* out);
int dashb_evaluate_compressed(dashb_circuit* c, const dashb_u128* const* in, uint64_t batch, dashb_u128* const* out);
/* dashb_evaluate_async with compressed labels on both sides (same rules; completes with dashb_evaluate_wait). */
int dashb_evaluate_compressed_async(dashb_circuit* c, const dashb_u128* const* in, uint64_t batch, dashb_u128* const* out);

/* Debug/parity hook: copies the output labels of layer `layer` (index into
 * the layers added, flatten not counted) of the last evaluation. */
int dashb_layer_outputs(dashb_circuit* c, int layer, int16_t* const* out);
/* decode_outputs (garbled_circuit_interface.h:422-467) on the device for
 * the last evaluation: values[batch][output_size], CRT-recombined and
 * mapped to the signed range; ok[batch] (optional) = 1 when every residue
 * matched its decoding information. */
int dashb_decode_outputs(dashb_circuit* c, int64_t* values, uint8_t* ok);

/* ---- output-channel sharding across GPUs (no counterpart in the reference,
 *      which is single-GPU) ----------------------
 * Every rank (one process per GPU) builds the same circuit, calls
 * dashb_set_shard(rank, world) before dashb_finalize and then owns, for each
 * linear layer, a contiguous block of output rows (dense rows / conv
 * filters) and, for the element-wise gadget layers that follow, the same
 * elements.  Weights and tables are replicated; only activations are
 * exchanged.  A rank evaluates a group of layers with
 * dashb_evaluate_layers(first, last) on its slices; before the next layer
 * that needs every input element (the next dense/conv, or the end) the
 * caller all-gathers the ranks' slices of the current planes -- plane j of
 * the current tensor is dashb_current_plane(): a device pointer to
 * n_elements rows of bytes_per_element bytes, rank r's slice of layer L's
 * output being elements dashb_layer_slice(L, r).  dash_b200/scheduler.py
 * does this with NCCL broadcasts over NVLink.  The exchange happens on the
 * digit planes as they are (no staging).  Sharded circuits run one lane. */
int dashb_set_shard(dashb_circuit* c, int rank, int world);
int dashb_evaluate_layers(dashb_circuit* c, int first, int last);
int dashb_current_plane(dashb_circuit* c, int j, void** dev_ptr, uint64_t* bytes_per_element,
                        uint64_t* n_elements);

/* The same exchange without NCCL: plain stores into peer memory over NVLink.
 * dashb_set_shard_lanes(c, n) (before dashb_finalize, n sub-batch lanes, 2 is
 * the useful number) gives every lane two gather buffers and a flag array;
 * every rank exports their CUDA IPC handles (dashb_shard_export: 3 x 64
 * bytes per lane), hands them to its peers by any host channel and imports
 * theirs (dashb_shard_import).  Then, per layer group and lane:
 * dashb_evaluate_layers_lane(lane, first, last) enqueues the layers on the
 * lane's stream and dashb_shard_exchange(lane, last) behind them one kernel
 * that stores this rank's slice into every rank's gather buffer, a signal and
 * a wait for all ranks' signals -- while one lane waits for its peers the
 * other lane computes.  Nothing synchronises with the host; a peer that never
 * signals makes the wait give up after a few seconds (dashb_shard_status).
 * All ranks must issue the same sequence of exchanges. */
int dashb_set_shard_lanes(dashb_circuit* c, int lanes);
int dashb_shard_lanes_used(const dashb_circuit* c);
/* the CUDA stream (cudaStream_t as void*) lane `lane` enqueues on, for a caller's own events */
void* dashb_lane_stream(const dashb_circuit* c, int lane);
int dashb_shard_export(dashb_circuit* c, int lane, void* handles_3x64);
int dashb_shard_import(dashb_circuit* c, int lane, int peer_rank, const void* handles_3x64);
int dashb_evaluate_layers_lane(dashb_circuit* c, int lane, int first, int last);
int dashb_shard_exchange(dashb_circuit* c, int lane, int layer);
/* all layer groups of all lanes with their exchanges (what a rank runs per
 * batch once the peers' buffers are imported); asynchronous */
int dashb_evaluate_sharded(dashb_circuit* c);
uint64_t dashb_shard_exchange_bytes(dashb_circuit* c, int lane, int layer);
int dashb_shard_status(dashb_circuit* c, int lane, uint32_t* timed_out_epoch);

/* Elements [e0, e1) of plane j of the current tensor as compressed labels in
 * DEVICE memory (dashb_u128 [batch][e1 - e0]) and back, enqueued on the
 * library stream: the sharded exchange then moves 16 bytes per label over
 * NVLink instead of 4*ceil(nc/4)*Bp/batch bytes of digits. */
int dashb_current_slice_compress(dashb_circuit* c, int j, uint64_t e0, uint64_t e1, void* dev_dst);
int dashb_current_slice_decompress(dashb_circuit* c, int j, uint64_t e0, uint64_t e1, const void* dev_src);
int dashb_layer_slice(const dashb_circuit* c, int layer, int rank, uint64_t* e0, uint64_t* e1);

/* ---- introspection ---------------------------------------------------- */
uint64_t dashb_input_size(const dashb_circuit* c);
uint64_t dashb_output_size(const dashb_circuit* c);
int dashb_num_layers(const dashb_circuit* c);
int dashb_crt_base_size(const dashb_circuit* c);
/* CRT modulus j (= modulus j of the circuit's input labels). */
int dashb_crt_modulus(const dashb_circuit* c, int j);
/* Modulus j of the circuit's output labels. */
int dashb_output_modulus(const dashb_circuit* c, int j);
/* Kernel launches issued by the last dashb_evaluate_device. */
uint64_t dashb_last_launch_count(const dashb_circuit* c);
/* Algorithmic bytes / AES blocks of one inference through layer `layer`
 * as defined in DESIGN.md (-1 = whole circuit). */
double dashb_algorithmic_bytes(const dashb_circuit* c, int layer);
double dashb_aes_blocks(const dashb_circuit* c, int layer);
double dashb_reference_aes_blocks(const dashb_circuit* c, int layer);
/* Device time (ms, CUDA events on dashb_stream()) each layer took in the
 * last evaluation run with profiling enabled; keys are layer indices. */
int dashb_set_layer_timing(dashb_circuit* c, int enabled);
int dashb_layer_time_ms(dashb_circuit* c, int layer, float* ms);
/* enabled = 2: the step runs as always (sub-batch lanes concurrent on their
 * own streams) with an event on the lane's stream before and after every
 * layer; after dashb_evaluate_device, dashb_layer_interval_ms gives the
 * interval [begin, end] of (lane, layer) in milliseconds since the step
 * began.  A layer's busy time in the step is the union of its lanes'
 * intervals.  (enabled = 1 serialises the lanes on one stream instead.) */
int dashb_lanes_used(const dashb_circuit* c);
int dashb_layer_interval_ms(dashb_circuit* c, int lane, int layer, float* begin_ms, float* end_ms);
/* Labels leaving layer `layer` per inference. */
uint64_t dashb_layer_size(const dashb_circuit* c, int layer);
/* Layer type of layer `layer` (DASHB_LAYER_*). */
int dashb_layer_type(const dashb_circuit* c, int layer);

#define DASHB_LAYER_DENSE 1
#define DASHB_LAYER_CONV2D 2
#define DASHB_LAYER_RELU 3
#define DASHB_LAYER_SIGN 4
#define DASHB_LAYER_RESCALE 5
#define DASHB_LAYER_RESCALE_LEGACY 6
#define DASHB_LAYER_PROJECTION 7
#define DASHB_LAYER_MAXPOOL2D 8
#define DASHB_LAYER_BASE_EXTENSION 9

/* ---- garbled-material files (no counterpart in the reference, which keeps
 *      a garbled circuit only in process memory; format in
 *      include/dash_gcm_format.h) ---------------------------------------- */
int dashb_circuit_load(dashb_circuit** out, const char* path);
/* ... with dashb_set_table_replicas(table_replicas) applied before the layers are added */
int dashb_circuit_load_replicated(dashb_circuit** out, const char* path, int table_replicas);
/* Parses a DGCM file on the host and checks every array's byte count against
 * what its layer's parameters imply (what dashb_circuit_load does before it
 * touches the device): DASHB_OK, DASHB_EIO (unreadable, truncated, a record
 * twice, a length that leaves the file) or DASHB_EINVAL (size mismatch).
 * Needs no device. */
int dashb_material_check(const char* path);

void dashb_circuit_destroy(dashb_circuit* c);

#ifdef __cplusplus
}
#endif
#endif /* DASH_B200_H */
