/* oracle_ops.c — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference ops on the hot path.
 * Pinned against the reference's ggml CPU backend (oracle/_ref) by tests/test_oracle_pin.py. */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

int orc_mul_mat_f64(int type, const void * w, int64_t w_row_bytes, int64_t M, int64_t K,
                    const float * x, int64_t N, float * dst) {
    float * row = (float *) malloc((size_t) K * sizeof(float));
    if (!row) return -1;
    for (int64_t m = 0; m < M; ++m) {
        if (orc_dequantize_row(type, (const uint8_t *) w + m * w_row_bytes, row, K) != 0) { free(row); return -1; }
        for (int64_t n = 0; n < N; ++n) {
            double acc = 0.0;
            for (int64_t k = 0; k < K; ++k) acc += (double) row[k] * (double) x[n * K + k];
            dst[n * M + m] = (float) acc;
        }
    }
    free(row);
    return 0;
}

/* integer value of weight k of a row (the value the CUDA reference feeds dp4a), plus scale terms */
int orc_mul_mat_q8_1(int type, const void * w, int64_t w_row_bytes, int64_t M, int64_t K,
                     const float * x, int64_t N, float * dst) {
    const int64_t nb = K / 32;
    uint8_t * xq = (uint8_t *) malloc((size_t) (nb * 36 * N));
    float * xd = (float *) malloc((size_t) (K * N) * sizeof(float));   /* dequantized q8_1 activations */
    float * row = (float *) malloc((size_t) K * sizeof(float));
    if (!xq || !xd || !row) { free(xq); free(xd); free(row); return -1; }
    for (int64_t n = 0; n < N; ++n) {
        orc_quantize_row_q8_1(x + n * K, xq + n * nb * 36, K);
        for (int64_t b = 0; b < nb; ++b) {
            const uint8_t * blk = xq + (n * nb + b) * 36;
            const float d = orc_fp16_to_fp32((uint16_t) (blk[0] | (blk[1] << 8)));
            for (int j = 0; j < 32; ++j) xd[n * K + b * 32 + j] = (float) (int8_t) blk[4 + j] * d;
        }
    }
    /* sum_k dequant(W)[k] * (d8*q8[k]) in f64 == the reference GEMV's value up to f32 rounding of
     * its scale products (vecdotq.cuh:103-122,231-243,295-314,461-483,580-600 all expand to this). */
    for (int64_t m = 0; m < M; ++m) {
        if (orc_dequantize_row(type, (const uint8_t *) w + m * w_row_bytes, row, K) != 0) { free(xq); free(xd); free(row); return -1; }
        for (int64_t n = 0; n < N; ++n) {
            double acc = 0.0;
            for (int64_t k = 0; k < K; ++k) acc += (double) row[k] * (double) xd[n * K + k];
            dst[n * M + m] = (float) acc;
        }
    }
    free(xq); free(xd); free(row);
    return 0;
}

double orc_nmse(const float * a, const float * b, size_t n) {   /* tests/test-backend-ops.cpp:276-289 */
    double mse_a_b = 0.0, mse_a_0 = 0.0;
    for (size_t i = 0; i < n; ++i) {
        const double d = (double) a[i] - (double) b[i];
        mse_a_b += d * d; mse_a_0 += (double) a[i] * (double) a[i];
    }
    return mse_a_b / mse_a_0;
}

void orc_rms_norm(const float * x, float * y, int64_t n, int64_t nrows, float eps) {
    for (int64_t r = 0; r < nrows; ++r) {
        double sum = 0.0;
        for (int64_t i = 0; i < n; ++i) sum += (double) x[r * n + i] * (double) x[r * n + i];
        const float mean = (float) (sum / (double) n);
        const float scale = 1.0f / sqrtf(mean + eps);
        for (int64_t i = 0; i < n; ++i) y[r * n + i] = x[r * n + i] * scale;
    }
}

void orc_glu(int op, const float * g, const float * u, float * out, int64_t n, float alpha, float limit) {
    for (int64_t i = 0; i < n; ++i) {
        float a = g[i], b = u[i], r;
        switch (op) {
            case 0: r = (a > 0 ? a : 0) * b; break;                                                             /* REGLU */
            case 1: r = 0.5f * a * (1.0f + tanhf(0.79788456080286535587989211986876f * a * (1.0f + 0.044715f * a * a))) * b; break; /* GEGLU */
            case 2: r = a / (1.0f + expf(-a)) * b; break;                                                       /* SWIGLU */
            case 3: a = a < limit ? a : limit; b = b < limit ? b : limit; b = b > -limit ? b : -limit;          /* SWIGLU_OAI */
                    r = a / (1.0f + expf(-a * alpha)) * (1.0f + b); break;
            case 4: r = 0.5f * a * (1.0f + erff(a * 0.70710678118654752440f)) * b; break;                       /* GEGLU_ERF */
            case 5: r = a / (1.0f + expf(-1.702f * a)) * b; break;                                              /* GEGLU_QUICK */
            default: r = b;
        }
        out[i] = r;
    }
}

void orc_rope(const float * x, float * y, int64_t hd, int64_t n_head, int64_t n_tok, const int32_t * pos,
              int n_dims, int mode, float freq_base, float freq_scale, const float * ff) {
    const float theta_scale = powf(freq_base, -2.0f / n_dims);
    for (int64_t t = 0; t < n_tok; ++t) {
        for (int64_t h = 0; h < n_head; ++h) {
            const float * s = x + (t * n_head + h) * hd; float * d = y + (t * n_head + h) * hd;
            float theta = (float) pos[t];
            for (int64_t i0 = 0; i0 < n_dims; i0 += 2) {
                const float f = ff ? ff[i0 / 2] : 1.0f;
                const float th = freq_scale * (theta / f);
                const float c = cosf(th), sn = sinf(th);
                const int64_t a = mode == 2 ? i0 / 2 : i0, b = mode == 2 ? i0 / 2 + n_dims / 2 : i0 + 1;
                d[a] = s[a] * c - s[b] * sn;
                d[b] = s[a] * sn + s[b] * c;
                theta *= theta_scale;
            }
            for (int64_t i = n_dims; i < hd; ++i) d[i] = s[i];
        }
    }
}

void orc_flash_attn(const float * q, const float * k, const float * v, const float * mask, const float * sinks,
                    float * out, int64_t D, int64_t n_q, int64_t H, int64_t H_kv, int64_t n_kv,
                    float scale, float max_bias, float logit_softcap) {
    const int64_t gqa = H / H_kv;
    const uint32_t n_head_log2 = 1u << (uint32_t) floor(log2((double) H));
    const float m0 = powf(2.0f, -(max_bias) / n_head_log2), m1 = powf(2.0f, -(max_bias / 2.0f) / n_head_log2);
    if (logit_softcap != 0.0f) scale /= logit_softcap;
    double * acc = (double *) malloc((size_t) D * sizeof(double));
    for (int64_t iq = 0; iq < n_q; ++iq) for (int64_t h = 0; h < H; ++h) {
        const float slope = max_bias > 0.0f ? (h < n_head_log2 ? powf(m0, h + 1) : powf(m1, 2 * (h - n_head_log2) + 1)) : 1.0f;
        const float * qv = q + (h * n_q + iq) * D;
        const int64_t hk = h / gqa;
        double M = -INFINITY, S = 0.0;
        for (int64_t d = 0; d < D; ++d) acc[d] = 0.0;
        for (int64_t ik = 0; ik < n_kv; ++ik) {
            const float mv = mask ? slope * mask[iq * n_kv + ik] : 0.0f;
            if (mv == -INFINITY) continue;
            const float * kv = k + (hk * n_kv + ik) * D;
            double s = 0.0;
            for (int64_t d = 0; d < D; ++d) s += (double) qv[d] * (double) kv[d];
            s *= scale;
            if (logit_softcap != 0.0f) s = logit_softcap * tanh(s);
            s += mv;
            const double Mold = M;
            double ms = 1.0, vs = 1.0;
            if (s > M) { M = s; ms = exp(Mold - M); for (int64_t d = 0; d < D; ++d) acc[d] *= ms; }
            else vs = exp(s - M);
            const float * vv = v + (hk * n_kv + ik) * D;
            for (int64_t d = 0; d < D; ++d) acc[d] += vs * (double) vv[d];
            S = S * ms + vs;
        }
        if (sinks) {
            const double s = sinks[h];
            double ms = 1.0, vs = 1.0;
            if (s > M) { ms = exp(M - s); M = s; for (int64_t d = 0; d < D; ++d) acc[d] *= ms; }
            else vs = exp(s - M);
            S = S * ms + vs;
        }
        for (int64_t d = 0; d < D; ++d) out[(iq * H + h) * D + d] = (float) (acc[d] / S);
    }
    free(acc);
}
