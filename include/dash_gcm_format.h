/*
 * dash_gcm_format.h -- "DGCM": a flat container for the garbled material of
 * one Dash circuit (what the evaluator holds after the offline phase).
 *
 * The reference keeps a garbled circuit only in process memory and has no
 * serialisation; this file defines the one used
 * here to carry material from the reference's garbler (integration/
 * dash_adapter.h) to the B200 evaluator, the CPU oracle and the test
 * fixtures.  Plain C99, header only, little endian, no arithmetic.
 *
 *   file    := "DASHGCM1" section*            (8-byte magic)
 *   section := u32 tag, u32 aux, u64 nbytes, payload padded to 16 bytes
 *
 * Section order: HEAD, ZERO, [UPSH], then per layer LAYR followed by its
 * ARRY sections in the fixed order listed at gcm_layer below, [DECI arrays],
 * [GINB/GINO arrays: the garbler-side input encoding, see gcm_circuit], ENDF.  All label arrays are int16 digits in the reference's
 * [label][digit] order; all ciphertext arrays are 16-byte little-endian
 * integers in the order the reference's garbler filled them (see
 * include/dash_b200.h for every table's layout and citation).
 */
#ifndef DASH_GCM_FORMAT_H
#define DASH_GCM_FORMAT_H

#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define GCM_TAG(a, b, c, d) \
    ((uint32_t)(a) | ((uint32_t)(b) << 8) | ((uint32_t)(c) << 16) | ((uint32_t)(d) << 24))
#define GCM_TAG_HEAD GCM_TAG('H', 'E', 'A', 'D')
#define GCM_TAG_ZERO GCM_TAG('Z', 'E', 'R', 'O')
#define GCM_TAG_UPSH GCM_TAG('U', 'P', 'S', 'H')
#define GCM_TAG_LAYR GCM_TAG('L', 'A', 'Y', 'R')
#define GCM_TAG_ARRY GCM_TAG('A', 'R', 'R', 'Y')
#define GCM_TAG_DECI GCM_TAG('D', 'E', 'C', 'I')
#define GCM_TAG_GINB GCM_TAG('G', 'I', 'N', 'B')
#define GCM_TAG_GINO GCM_TAG('G', 'I', 'N', 'O')
#define GCM_TAG_ENDF GCM_TAG('E', 'N', 'D', 'F')

#define GCM_MAX_PARAMS 32
#define GCM_MAX_ARRAYS 96

/* layer types (same numbering as DASHB_LAYER_* in dash_b200.h) */
#define GCM_DENSE 1
#define GCM_CONV2D 2
#define GCM_RELU 3
#define GCM_SIGN 4
#define GCM_RESCALE 5
#define GCM_RESCALE_LEGACY 6
#define GCM_PROJECTION 7
#define GCM_MAXPOOL2D 8
#define GCM_BASE_EXTENSION 9

typedef struct gcm_array {
    const void* data; /* not owned by views; owned (malloc) after gcm_load */
    uint64_t nbytes;
} gcm_array;

/*
 * p[0] = type, then
 *  DENSE          p[1]=M p[2]=K p[3]=channel_tf
 *                 arrays: W(int64 [M][K]), bias[k] (int16 [M][nc_j])
 *  CONV2D         p[1..8] = in_w in_h channels filters f_w f_h s_w s_h
 *                 arrays: W(int64), bias[k] (int16 [filters][nc_j])
 *  RELU           p[1]=n; arrays: approx cast1 cast2 sign garbler evaluator
 *  SIGN           p[1]=n; arrays: approx cast1 cast2 sign
 *  RESCALE        p[1]=n p[2]=nf p[3..3+nf)=factors
 *                 arrays: downshift[k], trans_mod, be_ciphers
 *  RESCALE_LEGACY p[1]=n p[2]=l
 *                 arrays: downshift[k], then per iteration
 *                         trans_mod approx cast1 cast2 sign
 *  PROJECTION     p[1]=n p[2..2+k)=in_moduli p[2+k..2+2k)=out_moduli
 *                 arrays: ciphers[k]
 *  BASE_EXTENSION p[1]=n p[2]=n_extra p[3..3+n_extra)=extra moduli (CRT moduli)
 *                 arrays: be_ciphers, then per extra modulus the labels the
 *                 reference layer starts that residue from (int16 [n][nc])
 *  MAXPOOL2D      p[1..7] = in_w in_h channels k_w k_h s_w s_h
 *                 arrays: approx cast1 cast2 sign garbler evaluator of the
 *                 outputs*k_w*k_h ReLU gadget instances (see dashb_add_maxpool2d)
 */
typedef struct gcm_layer {
    int64_t p[GCM_MAX_PARAMS];
    int n_arrays;
    gcm_array a[GCM_MAX_ARRAYS];
} gcm_layer;

typedef struct gcm_circuit {
    int k, n_mrs, max_modulus;
    uint64_t input_size;
    int32_t crt_base[32];
    int32_t mrs_base[32];
    gcm_array zero;    /* concatenated int16[nc(q)] for q = 2..max_modulus */
    gcm_array upshift; /* concatenated int16[nc(p_j)], nbytes 0 if absent   */
    int n_layers;
    gcm_layer* layers;
    int has_dec;
    gcm_array dec[32]; /* per CRT modulus u128 [p_j][output_size]           */
    /* Garbler-side input encoding (garble_inputs,
     * garbled_circuit_interface.h:325-343); optional -- an evaluator-only
     * deployment never holds it. */
    int has_ginp;
    gcm_array in_base[32];   /* per CRT modulus int16 [input_size][nc(p_j)] */
    gcm_array in_offset[32]; /* per CRT modulus int16 [nc(p_j)]: R_{p_j}    */
    int owns_data;     /* set by gcm_load: arrays are malloc'd              */
} gcm_circuit;

/* nc(q) exactly as LabelTensor::get_nr_comps (garbling/label_tensor.h:
 * 1353-1357): floor(128 / log2(q)) evaluated in double. */
static inline int gcm_nr_comps(int q) {
    return (int)floor(128 / log2((double)q));
}

static inline int gcm__wr(FILE* f, uint32_t tag, uint32_t aux, const void* p, uint64_t n) {
    static const char pad[16] = {0};
    if (fwrite(&tag, 4, 1, f) != 1 || fwrite(&aux, 4, 1, f) != 1 || fwrite(&n, 8, 1, f) != 1)
        return -1;
    if (n && fwrite(p, 1, n, f) != n) return -1;
    uint64_t r = (16 - (n & 15)) & 15;
    if (r && fwrite(pad, 1, r, f) != r) return -1;
    return 0;
}

static inline int gcm_save(const gcm_circuit* c, const char* path) {
    FILE* f = fopen(path, "wb");
    if (!f) return -1;
    int rc = 0;
    if (fwrite("DASHGCM1", 1, 8, f) != 8) rc = -1;
    int64_t head[8 + 64];
    memset(head, 0, sizeof(head));
    head[0] = c->k;
    head[1] = c->n_mrs;
    head[2] = c->max_modulus;
    head[3] = (int64_t)c->input_size;
    head[4] = c->n_layers;
    head[5] = c->upshift.nbytes ? 1 : 0;
    head[6] = c->has_dec;
    for (int i = 0; i < c->k; ++i) head[8 + i] = c->crt_base[i];
    for (int i = 0; i < c->n_mrs; ++i) head[8 + 32 + i] = c->mrs_base[i];
    rc |= gcm__wr(f, GCM_TAG_HEAD, 0, head, sizeof(head));
    rc |= gcm__wr(f, GCM_TAG_ZERO, 0, c->zero.data, c->zero.nbytes);
    if (c->upshift.nbytes) rc |= gcm__wr(f, GCM_TAG_UPSH, 0, c->upshift.data, c->upshift.nbytes);
    for (int l = 0; l < c->n_layers && !rc; ++l) {
        const gcm_layer* L = &c->layers[l];
        rc |= gcm__wr(f, GCM_TAG_LAYR, (uint32_t)L->n_arrays, L->p, sizeof(L->p));
        for (int a = 0; a < L->n_arrays; ++a)
            rc |= gcm__wr(f, GCM_TAG_ARRY, (uint32_t)a, L->a[a].data, L->a[a].nbytes);
    }
    if (c->has_dec)
        for (int j = 0; j < c->k; ++j)
            rc |= gcm__wr(f, GCM_TAG_DECI, (uint32_t)j, c->dec[j].data, c->dec[j].nbytes);
    if (c->has_ginp)
        for (int j = 0; j < c->k; ++j) {
            rc |= gcm__wr(f, GCM_TAG_GINB, (uint32_t)j, c->in_base[j].data, c->in_base[j].nbytes);
            rc |= gcm__wr(f, GCM_TAG_GINO, (uint32_t)j, c->in_offset[j].data, c->in_offset[j].nbytes);
        }
    rc |= gcm__wr(f, GCM_TAG_ENDF, 0, NULL, 0);
    if (fclose(f)) rc = -1;
    return rc;
}

/* Reads one record.  `left` = bytes of the file after the record header that
 * is about to be read; a length that does not fit (or whose padded size
 * wraps) is rejected before anything is allocated. */
static inline int gcm__rd(FILE* f, uint64_t* left, uint32_t* tag, uint32_t* aux, void** p, uint64_t* n) {
    *p = NULL;
    if (*left < 16) return -1;
    if (fread(tag, 4, 1, f) != 1 || fread(aux, 4, 1, f) != 1 || fread(n, 8, 1, f) != 1) return -1;
    *left -= 16;
    if (*n > ((uint64_t)1 << 40) || *n > *left) return -1;
    const uint64_t padded = (*n + 15) & ~(uint64_t)15;
    if (padded < *n || padded > *left) return -1;
    if (padded) {
        *p = malloc(padded);
        if (!*p) return -1;
        if (fread(*p, 1, padded, f) != padded) {
            free(*p);
            *p = NULL;
            return -1;
        }
        *left -= padded;
    }
    return 0;
}

static inline void gcm_free(gcm_circuit* c) {
    if (!c) return;
    if (c->owns_data) {
        free((void*)c->zero.data);
        free((void*)c->upshift.data);
        for (int l = 0; l < c->n_layers; ++l)
            for (int a = 0; a < c->layers[l].n_arrays; ++a) free((void*)c->layers[l].a[a].data);
        for (int j = 0; j < 32; ++j) {
            free((void*)c->dec[j].data);
            free((void*)c->in_base[j].data);
            free((void*)c->in_offset[j].data);
        }
    }
    free(c->layers);
    memset(c, 0, sizeof(*c));
}

/* Returns 0 on success; on failure the circuit is left zeroed. */
static inline int gcm_load(gcm_circuit* c, const char* path) {
    memset(c, 0, sizeof(*c));
    FILE* f = fopen(path, "rb");
    if (!f) return -1;
    char magic[8];
    if (fread(magic, 1, 8, f) != 8 || memcmp(magic, "DASHGCM1", 8) != 0) {
        fclose(f);
        return -1;
    }
    c->owns_data = 1;
    uint64_t left = 0;
    if (fseek(f, 0, SEEK_END) == 0) {
        const long end = ftell(f);
        if (end >= 8 && fseek(f, 8, SEEK_SET) == 0) left = (uint64_t)end - 8;
    }
    int cur = -1, rc = 0, done = 0, seen_head = 0;
    while (!rc && !done) {
        uint32_t tag, aux;
        uint64_t n;
        void* p;
        if (gcm__rd(f, &left, &tag, &aux, &p, &n)) {
            rc = -1;
            break;
        }
        /* every record but the header needs the header first; nothing may come twice */
        if ((tag != GCM_TAG_HEAD) != (seen_head != 0)) {
            free(p);
            rc = -1;
            break;
        }
        if (tag == GCM_TAG_HEAD) {
            seen_head = 1;
            if (n < (8 + 64) * 8) {
                rc = -1;
                free(p);
                break;
            }
            const int64_t* h = (const int64_t*)p;
            c->k = (int)h[0];
            c->n_mrs = (int)h[1];
            c->max_modulus = (int)h[2];
            c->input_size = (uint64_t)h[3];
            c->n_layers = (int)h[4];
            c->has_dec = (int)h[6];
            if (c->k < 1 || c->k > 32 || c->n_mrs < 0 || c->n_mrs > 32 || c->n_layers < 0) rc = -1;
            else {
                for (int i = 0; i < c->k; ++i) c->crt_base[i] = (int32_t)h[8 + i];
                for (int i = 0; i < c->n_mrs; ++i) c->mrs_base[i] = (int32_t)h[8 + 32 + i];
                c->layers = (gcm_layer*)calloc((size_t)(c->n_layers ? c->n_layers : 1), sizeof(gcm_layer));
                if (!c->layers) rc = -1;
            }
            free(p);
        } else if (tag == GCM_TAG_ZERO || tag == GCM_TAG_UPSH) {
            gcm_array* a = tag == GCM_TAG_ZERO ? &c->zero : &c->upshift;
            if (a->data) {
                rc = -1;
                free(p);
            } else {
                a->data = p;
                a->nbytes = n;
            }
        } else if (tag == GCM_TAG_LAYR) {
            ++cur;
            if (!c->layers || cur >= c->n_layers || n != sizeof(c->layers[0].p) || aux > GCM_MAX_ARRAYS) rc = -1;
            else {
                memcpy(c->layers[cur].p, p, n);
                c->layers[cur].n_arrays = (int)aux;
            }
            free(p);
        } else if (tag == GCM_TAG_ARRY) {
            if (cur < 0 || cur >= c->n_layers || (int)aux >= c->layers[cur].n_arrays || c->layers[cur].a[aux].data ||
                c->layers[cur].a[aux].nbytes) {
                rc = -1;
                free(p);
            } else {
                c->layers[cur].a[aux].data = p;
                c->layers[cur].a[aux].nbytes = n;
            }
        } else if (tag == GCM_TAG_DECI) {
            if (aux >= 32 || c->dec[aux].data) {
                rc = -1;
                free(p);
            } else {
                c->dec[aux].data = p;
                c->dec[aux].nbytes = n;
            }
        } else if (tag == GCM_TAG_GINB || tag == GCM_TAG_GINO) {
            gcm_array* a = aux < 32 ? (tag == GCM_TAG_GINB ? &c->in_base[aux] : &c->in_offset[aux]) : NULL;
            if (!a || a->data) {
                rc = -1;
                free(p);
            } else {
                a->data = p;
                a->nbytes = n;
                c->has_ginp = 1;
            }
        } else if (tag == GCM_TAG_ENDF) {
            free(p);
            done = 1;
        } else {
            free(p);
            rc = -1;
        }
    }
    fclose(f);
    if (!rc && cur + 1 != c->n_layers) rc = -1;
    if (rc) gcm_free(c);
    return rc;
}

/* Offset (in int16 units) of the zero label of modulus q inside c->zero. */
static inline uint64_t gcm_zero_offset(int q) {
    uint64_t off = 0;
    for (int m = 2; m < q; ++m) off += (uint64_t)gcm_nr_comps(m);
    return off;
}

#endif /* DASH_GCM_FORMAT_H */
