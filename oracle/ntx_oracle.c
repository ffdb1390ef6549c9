/* CPU oracle (plain C) for the sampling and compositing parts of the NeRF-Tex hot path.
 *
 * TEST INFRASTRUCTURE ONLY: linked/loaded only by tests/, __graft_entry__.smoke() and
 * bench.py's CPU-baseline legs.  The product library (libnerftex_b200.so) never calls it.
 *
 * PARITY UNPINNED: the reference arithmetic lives in tensorflow-gpu==2.4.1, absent from this
 * image; the reference has no tests or golden vectors.  This file is an independent second
 * restatement that must agree bit for bit with oracle/nerftex_oracle.py (same defined
 * left-to-right fp32 order, no FMA contraction: build with -ffp-contract=off).
 *
 * Citations are relative to /root/reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define NTXO_API __attribute__((visibility("default")))

/* tf.linspace(0,1,n)[i]: endpoints exact, interior delta*i with delta = fl(1/(n-1)). */
static float linspace01(int i, int n) {
    if (n == 1 || i == 0) return 0.0f;
    if (i == n - 1) return 1.0f;
    float delta = 1.0f / (float)(n - 1);
    return delta * (float)i;
}

/* network/renderer.py:101-111.  t[R,2], z_rand[R,S] (may be NULL when !perturb) -> z[R,S]. */
NTXO_API void ntxo_stratified_z(const float* t, const float* z_rand, int R, int S, int perturb, float* z) {
    float* base = (float*)malloc(sizeof(float) * (size_t)S);
    for (int r = 0; r < R; ++r) {
        float near = t[2 * r], far = t[2 * r + 1];
        for (int i = 0; i < S; ++i) {
            float tv = linspace01(i, S);
            float a = near * (1.0f - tv);
            float b = far * tv;
            base[i] = a + b;
        }
        for (int i = 0; i < S; ++i) {
            if (!perturb) { z[(size_t)r * S + i] = base[i]; continue; }
            float lower = (i == 0) ? base[0] : 0.5f * (base[i] + base[i - 1]);
            float upper = (i == S - 1) ? base[S - 1] : 0.5f * (base[i + 1] + base[i]);
            float d = upper - lower;
            float m = d * z_rand[(size_t)r * S + i];
            z[(size_t)r * S + i] = lower + m;
        }
    }
    free(base);
}

static float color_act(float c, int map_exr) {
    /* network/renderer.py:182-187 */
    if (map_exr) return c > 0.0f ? (float)((double)c + 1.0) : (float)exp((double)c);
    return (float)(1.0 / (1.0 + exp(-(double)c)));
}

/* network/renderer.py:170-213.  color[R,S,3], sigma[R,S], z[R,S], rays_d[R,3], noise[R,S] or NULL.
 * out: cmap[R,3], amap[R], weights[R,S], depth[R] (depth may be NULL). */
NTXO_API void ntxo_composite(const float* color, const float* sigma, const float* z, const float* rays_d,
                             const float* noise, int R, int S, int map_exr, int composite_bkgd,
                             const float* bkgd, float* cmap, float* amap, float* weights, float* depth) {
    for (int r = 0; r < R; ++r) {
        const float* d = rays_d + 3 * (size_t)r;
        float n2 = 0.0f;
        n2 = n2 + d[0] * d[0]; n2 = n2 + d[1] * d[1]; n2 = n2 + d[2] * d[2];
        float norm = sqrtf(n2);
        float T = 1.0f, c0 = 0.0f, c1 = 0.0f, c2 = 0.0f, acc = 0.0f, dep = 0.0f;
        for (int i = 0; i < S; ++i) {
            size_t k = (size_t)r * S + i;
            float dist;
            if (S == 1) dist = 0.0f; /* unreachable in the reference (needs S>=2) */
            else if (i < S - 1) dist = z[k + 1] - z[k];
            else dist = z[k] - z[k - 1];                       /* :177 last dist duplicated */
            dist = dist * norm;                                 /* :180 */
            float s = sigma[k];
            if (noise) s = s + noise[k];
            s = s > 0.0f ? s : 0.0f;
            float x = s * dist;
            float a = 1.0f - (float)exp(-(double)x);            /* :195 */
            float w = a * T;                                    /* :198 */
            float om = 1.0f - a;
            om = om + 1e-10f;
            T = T * om;
            weights[k] = w;
            float p0 = w * color_act(color[3 * k + 0], map_exr);
            float p1 = w * color_act(color[3 * k + 1], map_exr);
            float p2 = w * color_act(color[3 * k + 2], map_exr);
            c0 = c0 + p0; c1 = c1 + p1; c2 = c2 + p2;       /* :201 */
            float pz = w * z[k];
            dep = dep + pz;                                     /* :204 */
            acc = acc + w;                                      /* :207 */
        }
        if (composite_bkgd) {                                   /* :210-211 */
            float om = 1.0f - acc;
            float b0 = om * bkgd[0], b1 = om * bkgd[1], b2 = om * bkgd[2];
            c0 = c0 + b0; c1 = c1 + b1; c2 = c2 + b2;
        }
        cmap[3 * (size_t)r + 0] = c0; cmap[3 * (size_t)r + 1] = c1; cmap[3 * (size_t)r + 2] = c2;
        amap[r] = acc;
        if (depth) depth[r] = dep;
    }
}

/* network/renderer.py:589-618.  bins[R,C], weights[R,C-1], u[R,N] or NULL when det.
 * out: samples[R,N], inds[R,N] (may be NULL). */
NTXO_API void ntxo_sample_pdf(const float* bins, const float* weights, const float* u_in, int R, int C, int N,
                              int det, float* samples, int32_t* inds_out) {
    int W = C - 1;
    float* cdf = (float*)malloc(sizeof(float) * (size_t)C);
    for (int r = 0; r < R; ++r) {
        const float* w = weights + (size_t)r * W;
        const float* b = bins + (size_t)r * C;
        float total = 0.0f;
        for (int i = 0; i < W; ++i) { float wi = w[i] + 1e-5f; total = total + wi; }   /* :593-594 */
        cdf[0] = 0.0f;
        float run = 0.0f;
        for (int i = 0; i < W; ++i) {
            float wi = w[i] + 1e-5f;
            float p = wi / total;
            run = run + p;                                                            /* :595 */
            cdf[i + 1] = run;
        }
        for (int j = 0; j < N; ++j) {
            float u = det ? linspace01(j, N) : u_in[(size_t)r * N + j];               /* :599-603 */
            int lo = 0, hi = C;                                                       /* upper_bound, :606 */
            while (lo < hi) { int mid = (lo + hi) >> 1; if (cdf[mid] <= u) lo = mid + 1; else hi = mid; }
            int ind = lo;
            int below = ind - 1 > 0 ? ind - 1 : 0;                                    /* :607 */
            int above = ind < C - 1 ? ind : C - 1;                                    /* :608 */
            float denom = cdf[above] - cdf[below];                                    /* :613 */
            if (denom < 1e-5f) denom = 1.0f;                                          /* :614 */
            float tt = (u - cdf[below]) / denom;                                      /* :615 */
            float span = b[above] - b[below];
            float m = tt * span;
            samples[(size_t)r * N + j] = b[below] + m;                                /* :616 */
            if (inds_out) inds_out[(size_t)r * N + j] = ind;
        }
    }
    free(cdf);
}

static int cmp_float(const void* a, const void* b) {
    float x = *(const float*)a, y = *(const float*)b;
    return (x > y) - (x < y);
}

/* network/renderer.py:127-130: z_mid, sample_pdf(z_mid, weights[1:-1], N, det=perturb), sort(concat). */
NTXO_API void ntxo_hierarchical_z(const float* z, const float* weights, const float* u_in, int R, int S, int N,
                                  int det, float* z_out, float* z_samples, int32_t* inds_out) {
    float* mid = (float*)malloc(sizeof(float) * (size_t)(S - 1));
    for (int r = 0; r < R; ++r) {
        const float* zr = z + (size_t)r * S;
        for (int i = 0; i < S - 1; ++i) mid[i] = 0.5f * (zr[i + 1] + zr[i]);
        ntxo_sample_pdf(mid, weights + (size_t)r * S + 1, u_in ? u_in + (size_t)r * N : NULL, 1, S - 1, N, det,
                        z_samples + (size_t)r * N, inds_out ? inds_out + (size_t)r * N : NULL);
        float* o = z_out + (size_t)r * (S + N);
        memcpy(o, zr, sizeof(float) * (size_t)S);
        memcpy(o + S, z_samples + (size_t)r * N, sizeof(float) * (size_t)N);
        qsort(o, (size_t)(S + N), sizeof(float), cmp_float);
    }
    free(mid);
}
