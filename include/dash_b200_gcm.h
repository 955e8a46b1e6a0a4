/*
 * dash_b200_gcm.h -- builds a dashb_circuit from a gcm_circuit (a loaded
 * DGCM file or an in-memory view of a reference GarbledCircuit) by replaying
 * it through the C ABI's dashb_add_* entry points.  Header-only C99/C++.
 */
#ifndef DASH_B200_GCM_H
#define DASH_B200_GCM_H

#include "dash_b200.h"
#include "dash_gcm_format.h"

static inline int dashb_circuit_from_gcm(const gcm_circuit* g, dashb_circuit** out) {
    const int k = g->k;
    if (k < 1 || k > 32 || g->max_modulus < 2) return DASHB_EINVAL;
    const int nz = g->max_modulus - 1;
    const int16_t** zero = (const int16_t**)malloc(sizeof(void*) * (size_t)nz);
    if (!zero) return DASHB_ENOMEM;
    {
        const int16_t* z = (const int16_t*)g->zero.data;
        uint64_t off = 0;
        for (int q = 2; q <= g->max_modulus; ++q) {
            zero[q - 2] = z + off;
            off += (uint64_t)gcm_nr_comps(q);
        }
        if (off * 2 != g->zero.nbytes) {
            free(zero);
            return DASHB_EINVAL;
        }
    }
    const int16_t* up[32];
    if (g->upshift.nbytes) {
        const int16_t* u = (const int16_t*)g->upshift.data;
        uint64_t off = 0;
        for (int j = 0; j < k; ++j) {
            up[j] = u + off;
            off += (uint64_t)gcm_nr_comps(g->crt_base[j]);
        }
    }
    int rc = dashb_circuit_create(out, g->crt_base, k, g->mrs_base, g->n_mrs, g->max_modulus, zero,
                                  g->upshift.nbytes ? up : NULL, g->input_size);
    free(zero);
    if (rc != DASHB_OK) return rc;
    dashb_circuit* c = *out;
    for (int l = 0; l < g->n_layers && rc == DASHB_OK; ++l) {
        const gcm_layer* L = &g->layers[l];
        const void* ptr[GCM_MAX_ARRAYS];
        for (int a = 0; a < L->n_arrays; ++a) ptr[a] = L->a[a].data;
        switch ((int)L->p[0]) {
            case GCM_DENSE:
                if (L->n_arrays != 1 + k) rc = DASHB_EINVAL;
                else
                    rc = dashb_add_dense(c, (uint64_t)L->p[1], (uint64_t)L->p[2], (int)L->p[3],
                                         (const int64_t*)ptr[0], (const int16_t* const*)&ptr[1]);
                break;
            case GCM_CONV2D:
                if (L->n_arrays != 1 + k) rc = DASHB_EINVAL;
                else
                    rc = dashb_add_conv2d(c, (uint64_t)L->p[1], (uint64_t)L->p[2], (uint64_t)L->p[3],
                                          (uint64_t)L->p[4], (uint64_t)L->p[5], (uint64_t)L->p[6],
                                          (uint64_t)L->p[7], (uint64_t)L->p[8], (const int64_t*)ptr[0],
                                          (const int16_t* const*)&ptr[1]);
                break;
            case GCM_RELU:
                if (L->n_arrays != 6) rc = DASHB_EINVAL;
                else
                    rc = dashb_add_relu(c, (uint64_t)L->p[1], (const dashb_u128*)ptr[0],
                                        (const dashb_u128*)ptr[1], (const dashb_u128*)ptr[2],
                                        (const dashb_u128*)ptr[3], (const dashb_u128*)ptr[4],
                                        (const dashb_u128*)ptr[5]);
                break;
            case GCM_SIGN:
                if (L->n_arrays != 4) rc = DASHB_EINVAL;
                else
                    rc = dashb_add_sign(c, (uint64_t)L->p[1], (const dashb_u128*)ptr[0],
                                        (const dashb_u128*)ptr[1], (const dashb_u128*)ptr[2],
                                        (const dashb_u128*)ptr[3]);
                break;
            case GCM_RESCALE: {
                int32_t factors[16];
                int nf = (int)L->p[2];
                if (nf < 1 || nf > 16 || L->n_arrays != k + 2) {
                    rc = DASHB_EINVAL;
                    break;
                }
                for (int i = 0; i < nf; ++i) factors[i] = (int32_t)L->p[3 + i];
                rc = dashb_add_rescale(c, (uint64_t)L->p[1], factors, nf, (const int16_t* const*)&ptr[0],
                                       (const dashb_u128*)ptr[k], (const dashb_u128*)ptr[k + 1]);
                break;
            }
            case GCM_RESCALE_LEGACY: {
                int l_it = (int)L->p[2];
                if (l_it < 1 || l_it > 16 || L->n_arrays != k + 5 * l_it) {
                    rc = DASHB_EINVAL;
                    break;
                }
                const dashb_u128 *tm[16], *ap[16], *c1[16], *c2[16], *sg[16];
                for (int it = 0; it < l_it; ++it) {
                    tm[it] = (const dashb_u128*)ptr[k + 5 * it + 0];
                    ap[it] = (const dashb_u128*)ptr[k + 5 * it + 1];
                    c1[it] = (const dashb_u128*)ptr[k + 5 * it + 2];
                    c2[it] = (const dashb_u128*)ptr[k + 5 * it + 3];
                    sg[it] = (const dashb_u128*)ptr[k + 5 * it + 4];
                }
                rc = dashb_add_rescale_legacy(c, (uint64_t)L->p[1], l_it, (const int16_t* const*)&ptr[0], tm, ap,
                                              c1, c2, sg);
                break;
            }
            case GCM_PROJECTION: {
                int32_t im[32], om[32];
                if (L->n_arrays != k) {
                    rc = DASHB_EINVAL;
                    break;
                }
                for (int j = 0; j < k; ++j) {
                    im[j] = (int32_t)L->p[2 + j];
                    om[j] = (int32_t)L->p[2 + k + j];
                }
                rc = dashb_add_projection(c, (uint64_t)L->p[1], im, om, (const dashb_u128* const*)&ptr[0]);
                break;
            }
            default:
                rc = DASHB_EINVAL;
        }
    }
    if (rc == DASHB_OK && g->has_dec) {
        const dashb_u128* dec[32];
        for (int j = 0; j < k; ++j) dec[j] = (const dashb_u128*)g->dec[j].data;
        rc = dashb_set_decoding_info(c, dec);
    }
    if (rc == DASHB_OK && g->has_ginp) {
        const int16_t *base[32], *off[32];
        for (int j = 0; j < k; ++j) {
            base[j] = (const int16_t*)g->in_base[j].data;
            off[j] = (const int16_t*)g->in_offset[j].data;
        }
        rc = dashb_set_input_encoding(c, base, off);
    }
    if (rc != DASHB_OK) {
        dashb_circuit_destroy(c);
        *out = NULL;
    }
    return rc;
}

#endif /* DASH_B200_GCM_H */
