/*
 * dash_b200_gcm.h -- builds a dashb_circuit from a gcm_circuit (a loaded
 * DGCM file or an in-memory view of a reference GarbledCircuit) by replaying
 * it through the C ABI's dashb_add_* entry points.  Header-only C99/C++.
 */
#ifndef DASH_B200_GCM_H
#define DASH_B200_GCM_H

#include "dash_b200.h"
#include "dash_gcm_format.h"

/* ---- size checks: every array of a DGCM file against the byte count its
 *      layer's parameters imply, before any pointer reaches dashb_add_* (a
 *      truncated or hand-made file must not make the library read past a
 *      buffer).  Overflowing products count as a mismatch. ---- */
static inline uint64_t gcm__mul(uint64_t a, uint64_t b, int* bad) {
    uint64_t r;
    if (__builtin_mul_overflow(a, b, &r)) {
        *bad = 1;
        return 0;
    }
    return r;
}
static inline int gcm__is(const gcm_array* a, uint64_t want, int bad) {
    return !bad && a->nbytes == want && (want == 0 || a->data != NULL);
}

/* 1 when layer L's arrays have exactly the sizes its parameters imply.
 * n_in = labels entering the layer; *n_out = labels leaving it. */
static inline int gcm_layer_sizes_ok(const gcm_circuit* g, const gcm_layer* L, uint64_t n_in, uint64_t* n_out) {
    const int k = g->k;
    int bad = 0;
    uint64_t S = 0, Mp = 0, nc[32];
    for (int j = 0; j < k; ++j) {
        if (g->crt_base[j] < 2) return 0;
        S += (uint64_t)g->crt_base[j];
        nc[j] = (uint64_t)gcm_nr_comps(g->crt_base[j]);
    }
    for (int i = 0; i < g->n_mrs; ++i) {
        if (g->mrs_base[i] < 2) return 0;
        if (i) Mp += (uint64_t)g->mrs_base[i];
    }
    for (int i = 1; i <= 8; ++i)
        if (L->p[i] < 0) return 0;
    const uint64_t n = (uint64_t)L->p[1];
    *n_out = n_in;
    const int type = (int)L->p[0];
    /* the four SignGadget tables starting at array index a0, sign table for n_out moduli */
    #define GCM__SIGN_OK(a0, nout)                                                                              \
        (g->n_mrs >= 1 &&                                                                                       \
         gcm__is(&L->a[(a0) + 0], gcm__mul(gcm__mul(n, (uint64_t)g->n_mrs, &bad), gcm__mul(S, 16, &bad), &bad), bad) && \
         gcm__is(&L->a[(a0) + 1], gcm__mul(gcm__mul(n, (uint64_t)(k + 1), &bad), gcm__mul(Mp, 16, &bad), &bad), bad) && \
         gcm__is(&L->a[(a0) + 2], gcm__mul(gcm__mul(n, (uint64_t)(k + 1), &bad), gcm__mul(Mp, 16, &bad), &bad), bad) && \
         gcm__is(&L->a[(a0) + 3], gcm__mul(gcm__mul(n, (uint64_t)(nout), &bad), gcm__mul((uint64_t)g->mrs_base[0], 16, &bad), &bad), bad))
    switch (type) {
        case GCM_DENSE: {
            const uint64_t M = (uint64_t)L->p[1], K = (uint64_t)L->p[2];
            if (L->n_arrays != 1 + k || K != n_in) return 0;
            if (!gcm__is(&L->a[0], gcm__mul(gcm__mul(M, K, &bad), 8, &bad), bad)) return 0;
            for (int j = 0; j < k; ++j)
                if (!gcm__is(&L->a[1 + j], gcm__mul(gcm__mul(M, nc[j], &bad), 2, &bad), bad)) return 0;
            *n_out = M;
            return 1;
        }
        case GCM_CONV2D: {
            const uint64_t in_w = (uint64_t)L->p[1], in_h = (uint64_t)L->p[2], ch = (uint64_t)L->p[3], fl = (uint64_t)L->p[4],
                           fw = (uint64_t)L->p[5], fh = (uint64_t)L->p[6], sw = (uint64_t)L->p[7], sh = (uint64_t)L->p[8];
            if (L->n_arrays != 1 + k || !sw || !sh || !fw || !fh || fw > in_w || fh > in_h) return 0;
            if (gcm__mul(gcm__mul(in_w, in_h, &bad), ch, &bad) != n_in || bad) return 0;
            if (!gcm__is(&L->a[0], gcm__mul(gcm__mul(gcm__mul(fl, ch, &bad), gcm__mul(fw, fh, &bad), &bad), 8, &bad), bad)) return 0;
            for (int j = 0; j < k; ++j)
                if (!gcm__is(&L->a[1 + j], gcm__mul(gcm__mul(fl, nc[j], &bad), 2, &bad), bad)) return 0;
            *n_out = gcm__mul(gcm__mul((in_w - fw) / sw + 1, (in_h - fh) / sh + 1, &bad), fl, &bad);
            return !bad;
        }
        case GCM_RELU:
            if (L->n_arrays != 6 || n != n_in || !GCM__SIGN_OK(0, 1)) return 0;
            return gcm__is(&L->a[4], gcm__mul(gcm__mul(n, S, &bad), 16, &bad), bad) &&
                   gcm__is(&L->a[5], gcm__mul(gcm__mul(n, (uint64_t)k, &bad), 48, &bad), bad);
        case GCM_SIGN: return L->n_arrays == 4 && n == n_in && GCM__SIGN_OK(0, k);
        case GCM_MAXPOOL2D: {
            const uint64_t in_w = (uint64_t)L->p[1], in_h = (uint64_t)L->p[2], ch = (uint64_t)L->p[3], kw = (uint64_t)L->p[4],
                           kh = (uint64_t)L->p[5], sw = (uint64_t)L->p[6], sh = (uint64_t)L->p[7];
            if (L->n_arrays != 6 || !kw || !kh || !sw || !sh || kw > in_w || kh > in_h) return 0;
            if (gcm__mul(gcm__mul(in_w, in_h, &bad), ch, &bad) != n_in || bad) return 0;
            *n_out = gcm__mul(gcm__mul((in_w - kw) / sw + 1, (in_h - kh) / sh + 1, &bad), ch, &bad);
            {
                const uint64_t n = gcm__mul(*n_out, gcm__mul(kw, kh, &bad), &bad); /* gadget instances */
                if (!GCM__SIGN_OK(0, 1)) return 0;
                return gcm__is(&L->a[4], gcm__mul(gcm__mul(n, S, &bad), 16, &bad), bad) &&
                       gcm__is(&L->a[5], gcm__mul(gcm__mul(n, (uint64_t)k, &bad), 48, &bad), bad);
            }
        }
        case GCM_RESCALE: {
            const int nf = (int)L->p[2];
            if (nf < 1 || nf > 16 || nf >= k || L->n_arrays != k + 2 || n != n_in) return 0;
            for (int j = 0; j < k; ++j)
                if (!gcm__is(&L->a[j], nc[j] * 2, 0)) return 0;
            /* trans_mod: factor i serves the k-1-i moduli not yet ignored; base extension over the base
             * with the factors swapped to the end (dashb_add_rescale) */
            int perm[32], fidx[16];
            for (int j = 0; j < k; ++j) perm[j] = j;
            uint64_t tm = 0, be = 0;
            for (int i = 0; i < nf; ++i) {
                int f = -1;
                for (int j = 0; j < k; ++j)
                    if (g->crt_base[j] == (int32_t)L->p[3 + i] && f < 0) f = j;
                if (f < 0) return 0;
                fidx[i] = f;
                tm += gcm__mul(gcm__mul((uint64_t)(k - 1 - i), n, &bad), (uint64_t)L->p[3 + i], &bad);
            }
            for (int i = 1; i <= nf; ++i) {
                const int a = fidx[nf - i], b = k - i, t = perm[a];
                perm[a] = perm[b];
                perm[b] = t;
            }
            for (int i = 0; i < k - nf; ++i)
                be += gcm__mul(gcm__mul((uint64_t)(k - i - 1), n, &bad), (uint64_t)g->crt_base[perm[i]], &bad);
            return gcm__is(&L->a[k], gcm__mul(tm, 16, &bad), bad) && gcm__is(&L->a[k + 1], gcm__mul(be, 16, &bad), bad);
        }
        case GCM_BASE_EXTENSION: {
            const int nx = (int)L->p[2];
            if (nx < 1 || nx > 16 || nx >= k || L->n_arrays != 1 + nx || n != n_in) return 0;
            int perm[32], fidx[16];
            for (int j = 0; j < k; ++j) perm[j] = j;
            for (int i = 0; i < nx; ++i) {
                int f = -1;
                for (int j = 0; j < k; ++j)
                    if (g->crt_base[j] == (int32_t)L->p[3 + i] && f < 0) f = j;
                if (f < 0) return 0;
                fidx[i] = f;
                if (!gcm__is(&L->a[1 + i], gcm__mul(gcm__mul(n, nc[f], &bad), 2, &bad), bad)) return 0;
            }
            for (int i = 1; i <= nx; ++i) {
                const int a = fidx[nx - i], b = k - i, t = perm[a];
                perm[a] = perm[b];
                perm[b] = t;
            }
            uint64_t be = 0;
            for (int i = 0; i < k - nx; ++i)
                be += gcm__mul(gcm__mul((uint64_t)(k - i - 1), n, &bad), (uint64_t)g->crt_base[perm[i]], &bad);
            return gcm__is(&L->a[0], gcm__mul(be, 16, &bad), bad);
        }
        case GCM_RESCALE_LEGACY: {
            const int l_it = (int)L->p[2];
            if (l_it < 1 || l_it > 16 || L->n_arrays != k + 5 * l_it || n != n_in) return 0;
            for (int j = 0; j < k; ++j)
                if (!gcm__is(&L->a[j], nc[j] * 2, 0)) return 0;
            for (int it = 0; it < l_it; ++it) {
                if (!gcm__is(&L->a[k + 5 * it], gcm__mul(gcm__mul(n, (uint64_t)(k - 1), &bad), 32, &bad), bad)) return 0;
                if (!GCM__SIGN_OK(k + 5 * it + 1, 1)) return 0;
            }
            return 1;
        }
        case GCM_PROJECTION:
            if (L->n_arrays != k || n != n_in || 2 + 2 * k > GCM_MAX_PARAMS) return 0;
            for (int j = 0; j < k; ++j)
                if (L->p[2 + j] < 2 || !gcm__is(&L->a[j], gcm__mul(gcm__mul(n, (uint64_t)L->p[2 + j], &bad), 16, &bad), bad)) return 0;
            return 1;
        default: return 0;
    }
    #undef GCM__SIGN_OK
}

/* 1 when every array of the file has the size its parameters imply. */
static inline int gcm_sizes_ok(const gcm_circuit* g) {
    const int k = g->k;
    if (k < 1 || k > 32 || g->n_mrs < 0 || g->n_mrs > 32 || g->max_modulus < 2 || g->n_layers < 0 || !g->input_size) return 0;
    if (g->n_layers && !g->layers) return 0;
    uint64_t up = 0, cur = g->input_size;
    for (int j = 0; j < k; ++j) {
        if (g->crt_base[j] < 2) return 0;
        up += (uint64_t)gcm_nr_comps(g->crt_base[j]) * 2;
    }
    if (g->upshift.nbytes && !gcm__is(&g->upshift, up, 0)) return 0;
    for (int l = 0; l < g->n_layers; ++l) {
        uint64_t next = 0;
        if (g->layers[l].n_arrays < 0 || g->layers[l].n_arrays > GCM_MAX_ARRAYS) return 0;
        if (!gcm_layer_sizes_ok(g, &g->layers[l], cur, &next)) return 0;
        cur = next;
    }
    for (int j = 0; j < k; ++j) {
        int bad = 0;
        if (g->has_dec && !gcm__is(&g->dec[j], gcm__mul(gcm__mul((uint64_t)g->crt_base[j], cur, &bad), 16, &bad), bad)) return 0;
        if (g->has_ginp) {
            const uint64_t ncj = (uint64_t)gcm_nr_comps(g->crt_base[j]);
            if (!gcm__is(&g->in_base[j], gcm__mul(gcm__mul(g->input_size, ncj, &bad), 2, &bad), bad)) return 0;
            if (!gcm__is(&g->in_offset[j], ncj * 2, 0)) return 0;
        }
    }
    return 1;
}

static inline int dashb_circuit_from_gcm(const gcm_circuit* g, dashb_circuit** out) {
    const int k = g->k;
    if (k < 1 || k > 32 || g->max_modulus < 2) return DASHB_EINVAL;
    if (!gcm_sizes_ok(g)) return DASHB_EINVAL;
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
            case GCM_MAXPOOL2D:
                if (L->n_arrays != 6) rc = DASHB_EINVAL;
                else
                    rc = dashb_add_maxpool2d(c, (uint64_t)L->p[1], (uint64_t)L->p[2], (uint64_t)L->p[3],
                                             (uint64_t)L->p[4], (uint64_t)L->p[5], (uint64_t)L->p[6], (uint64_t)L->p[7],
                                             (const dashb_u128*)ptr[0], (const dashb_u128*)ptr[1],
                                             (const dashb_u128*)ptr[2], (const dashb_u128*)ptr[3],
                                             (const dashb_u128*)ptr[4], (const dashb_u128*)ptr[5]);
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
            case GCM_BASE_EXTENSION: {
                int32_t extra[16];
                int nx = (int)L->p[2];
                if (nx < 1 || nx > 16 || L->n_arrays != 1 + nx) {
                    rc = DASHB_EINVAL;
                    break;
                }
                for (int i = 0; i < nx; ++i) extra[i] = (int32_t)L->p[3 + i];
                rc = dashb_add_base_extension(c, (uint64_t)L->p[1], extra, nx, (const int16_t* const*)&ptr[1],
                                              (const dashb_u128*)ptr[0]);
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
