/*
 * dash_b200_garble.h -- the offline phase on the GPU (SURVEY.md section 8,
 * "next" row f3): garbles a plaintext circuit into a DGCM material file
 * (include/dash_gcm_format.h) that dashb_circuit_load, the CPU oracle and any
 * other consumer of the format can evaluate.
 *
 * The reference garbles on the host, gate by gate, with RDRAND labels
 * (GarbledCircuit::garble, garbling/garbled_circuit.h:122-176; the gadgets'
 * garble() methods cited at each layer type below) and calls the result "not
 * optimized" (README.md:70).  This entry point produces material of the same
 * structure -- every table in the reference's layout, see include/dash_b200.h
 * -- from a seed: labels are AES-128 outputs under a key derived from the
 * seed, so the same seed gives the same circuit.  It is NOT label-for-label
 * equal to what the reference would produce for the same circuit (the
 * reference's labels are hardware random numbers); correctness is defined by
 * evaluation: evaluate(garble(C), encode(x)) decodes to C(x), checked against
 * the pinned CPU oracle and against the reference-garbled circuit's own
 * decoded outputs (tests/test_gpu_garbler.py).
 *
 * Supported layers: dense, conv2d, ReLU (approximate sign gadget + mixed-
 * modulus half gates), sign, rescale with base extension, legacy rescale.
 * Not supported: standalone projection layers.
 */
#ifndef DASH_B200_GARBLE_H
#define DASH_B200_GARBLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/*
 * One layer of the plaintext circuit (circuit/layer/*.h).  p[] as in
 * gcm_layer (include/dash_gcm_format.h): p[0] = DASHB_LAYER_* type, then
 *   DENSE    p[1]=M p[2]=K p[3]=channel_tf      weights int64 [M][K], biases int64 [M]
 *   CONV2D   p[1..8]=in_w in_h channels filters f_w f_h s_w s_h
 *            weights in the reference's order (label_tensor.h:967-969), biases [filters]
 *   RELU     (n is the previous layer's output size)
 *   SIGN
 *   RESCALE  p[2]=nf p[3..3+nf)=factors (each a CRT modulus)
 *   RESCALE_LEGACY  p[2]=l (iterations of "halve + sign gadget"; needs crt_base[0] == 2)
 * Weights and biases are the quantised integers of Dense::get_q_weights /
 * get_q_biases (garbling/layer/garbled_dense.h:91-96); any representative
 * modulo the CRT modulus is accepted.
 */
typedef struct dashb_plain_layer {
    int64_t p[16];
    const int64_t* weights;
    const int64_t* biases;
} dashb_plain_layer;

typedef struct dashb_garble_stats {
    double kernel_ms;      /* device time of the gate-garbling kernels (CUDA events) */
    double total_ms;       /* wall time of the whole call, file write included */
    uint64_t gates;        /* projection gates garbled */
    uint64_t rows;         /* ciphertext rows written (= AES blocks hashed for tables) */
    uint64_t table_bytes;
} dashb_garble_stats;

/*
 * Garbles the circuit and writes the material, the decoding information and
 * the input encoding (GINB/GINO sections) to `path`.  `seed` is the AES key of
 * the PRF every label (the offsets R_q included) is drawn from: it must be
 * secret, fresh randomness in a deployment; a constant is for reproducible
 * tests only.  max_modulus is raised
 * to the smallest value that covers every modulus the gadgets use (as the
 * reference's constructors do, garbled_circuit_interface.h:128-135).  Needs
 * dashb_init() on the current device.  Returns DASHB_OK or a DASHB_E* code
 * (dashb_last_error() has the text).
 */
int dashb_garble_to_file(const int32_t* crt_base, int k, const int32_t* mrs_base, int n_mrs, int max_modulus,
                         uint64_t input_size, const dashb_plain_layer* layers, int n_layers, const uint8_t seed[16],
                         const char* path, dashb_garble_stats* stats);

/*
 * Same garbling, written as two files: `evaluator_path` holds what the
 * evaluating party needs and nothing else (tables, weights, the bias / zero /
 * shift labels); `garbler_path` holds what must stay with the garbler -- the
 * input base labels, the offset labels R_q (GINB/GINO sections) and the
 * decoding information -- as a DGCM file without layers.  Whoever holds the
 * second file can open every input and output label; the single-file form
 * above is for tests, benchmarks and single-party use only.  No counterpart
 * in the reference (it never serialises a circuit).
 */
int dashb_garble_to_files(const int32_t* crt_base, int k, const int32_t* mrs_base, int n_mrs, int max_modulus,
                          uint64_t input_size, const dashb_plain_layer* layers, int n_layers, const uint8_t seed[16],
                          const char* evaluator_path, const char* garbler_path, dashb_garble_stats* stats);

/*
 * Same, straight into an evaluator circuit on the current device: the
 * ciphertext tables never leave the GPU (device-to-device copies into the
 * circuit).  The circuit comes back un-finalized, with decoding information
 * and input encoding set: call dashb_finalize(*circuit, max_batch), then
 * dashb_garble_inputs / dashb_evaluate_device / dashb_decode_outputs.
 */
struct dashb_circuit;
int dashb_garble_to_circuit(const int32_t* crt_base, int k, const int32_t* mrs_base, int n_mrs, int max_modulus,
                            uint64_t input_size, const dashb_plain_layer* layers, int n_layers, const uint8_t seed[16],
                            struct dashb_circuit** circuit, dashb_garble_stats* stats);

#ifdef __cplusplus
}
#endif
#endif /* DASH_B200_GARBLE_H */
