/*
 * dash_oracle.h -- TEST INFRASTRUCTURE (the parity checker), not product.
 * C API of the plain-C restatement of Dash's online evaluator; see
 * dash_oracle.c for the reference citations of every function.
 */
#ifndef DASH_ORACLE_H
#define DASH_ORACLE_H
#include <stdint.h>

typedef struct oracle_circuit oracle_circuit;

/* primitives (label_tensor.h:715-742,775-780; cpu_aes_engine.h:23-40) */
void oracle_aes128_fixed_key(const uint8_t in[16], uint8_t out[16]);
void oracle_compress(int q, const int16_t* digits, uint64_t out2[2]);
void oracle_decompress(int q, const uint64_t in2[2], int16_t* digits);
void oracle_hash(const uint64_t in2[2], uint64_t out2[2]);
int oracle_nr_comps(int q);
int64_t oracle_mul_inv(uint64_t a, int b);

oracle_circuit* oracle_load(const char* path);
void oracle_free(oracle_circuit* c);
int oracle_k(const oracle_circuit* c);
int oracle_num_layers(const oracle_circuit* c);
uint64_t oracle_input_size(const oracle_circuit* c);
uint64_t oracle_layer_out_size(const oracle_circuit* c, int layer); /* -1 = last */
int oracle_layer_out_modulus(const oracle_circuit* c, int layer, int j);

/* Evaluates layers 0..upto (upto = -1: all) on one inference.
 * in[j] = int16 [input_size][nc(p_j)], out[j] = int16 [size][nc(q_j)] of
 * layer `upto`'s output.  Returns 0, or -1 on a malformed circuit. */
int oracle_evaluate(const oracle_circuit* c, const int16_t* const* in, int upto, int16_t* const* out,
                    int nr_threads);
/* decode_outputs (garbled_circuit_interface.h:422-467).  Returns the number
 * of residues that did not match any decoding hash (0 = decoded). */
int oracle_decode(const oracle_circuit* c, const int16_t* const* labels, int64_t* values);
/* AES blocks the reference evaluator computes for one inference. */
double oracle_aes_blocks(const oracle_circuit* c);

#endif
