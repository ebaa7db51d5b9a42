/* TEST INFRASTRUCTURE (oracle/): tier-2 oracle = FP64 C restatement of the reference hot path.
 *
 * Op-for-op restatement (same operation order; no implicit FMA contraction: build with
 * -ffp-contract=off; the only fused operations are the explicit fma() calls in dot3) of thiagoald/pathtracerpython:
 *   utils.py:98-147   intersect          -> pto_intersect_one
 *   utils.py:72-91    in_triangle        -> in_triangle
 *   utils.py:48-49    squared_dist       -> squared_dist
 *   utils.py:21-46    sample_bary_coords / pick_random_triangle / sample_random_pt
 *   utils.py:55-69    make_rays / make_screen_pts (numpy.linspace semantics)
 *   utils.py:150-161  make_image         -> pto_make_image
 *   main.py:23-73     compute_shadow_rays (incl. the leaked loop variable `obj`, main.py:42,70)
 *   main.py:76-80     compute_ambient_color
 *   main.py:83-122    intersect_objects  (objects in file order, light mesh last, first minimum)
 *   main.py:148-162   rotate
 *   main.py:186-280   sample loop / bounce loop / accumulation / mean
 * Randomness: the injected Philox stream of oracle/philox.py (same dimension map).
 *
 * PARITY PIN: the reference ships no tests or golden vectors; this file is pinned
 * against vectors produced by RUNNING the unmodified reference functions in the build container
 * (oracle/ref_driver.py -> tests/golden/, script tests/golden/make_golden.py), checked by
 * tests/test_oracle_golden.py.
 *
 * This is the checker and the timed CPU baseline only.  Nothing under pathtracerpython_b200/
 * links, imports or calls it.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    int32_t n_tri;          /* closest-hit set: object triangles then light triangles (main.py:91) */
    int32_t n_tri_obj;      /* occlusion set: object triangles only (main.py:42) */
    int32_t n_obj;          /* number of scene.objects; light triangles carry object index n_obj */
    int32_t n_light_tri;
    const double *verts;    /* [n_tri][3][3] v1,v2,v3 as parsed (scene_reader.py:83-86) */
    const double *normals;  /* [n_tri][3] stored normal = calc_normal (scene_reader.py:5-8) */
    const int32_t *tri_obj; /* [n_tri] object index */
    const double *mat;      /* [n_obj][8] red green blue ka kd ks kt n (scene_reader.py:175-182) */
    const double *light_areas; /* [n_light_tri] (scene_reader.py:88) */
    double eye[3];
    double ortho[4];        /* x0 y0 x1 y1 */
    int32_t width, height;
    double ambient;
    double light_color[3];
} pto_scene;

#define ZERO 1e-5   /* main.py:20, utils.py:18 */
#define TAU 6.28    /* main.py:19 */

/* ---------------- Philox4x32-10 (oracle/philox.py) ---------------- */
static void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

void pto_philox(const uint32_t *ctr, const uint32_t *key, uint32_t *out) { philox4x32_10(ctr, key, out); }

static double u01(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, uint32_t dim) {
    uint32_t ctr[4] = {pixel, sample, bounce, dim >> 2}, key[2] = {seed, 0}, o[4];
    philox4x32_10(ctr, key, o);
    return (double)(o[dim & 3] >> 8) * 0x1p-24;
}
/* random.uniform(a, b) = a + (b - a) * random() */
static double uniform_ab(double a, double b, double u) { return a + (b - a) * u; }

/* ---------------- small vector helpers in numpy's operation order ---------------- */
/* np.dot / np.linalg.norm on 3-vectors go through BLAS ddot, which on the machine that produced the
 * goldens evaluates fma(a2,b2, fma(a1,b1, a0*b0)) (probed: 100% bit-identical over 1e5 random pairs;
 * the plain sequential sum matches only 67%).  Restated with explicit fma(); everything else keeps
 * one rounding per operation. */
static double dot3(const double *a, const double *b) { return fma(a[2], b[2], fma(a[1], b[1], a[0] * b[0])); }
static double norm3(const double *a) { return sqrt(dot3(a, a)); }            /* np.linalg.norm */
static void cross3(const double *a, const double *b, double *c) {            /* np.cross */
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}
static void sub3(const double *a, const double *b, double *c) { c[0] = a[0] - b[0]; c[1] = a[1] - b[1]; c[2] = a[2] - b[2]; }
static void div3(double *a, double s) { a[0] /= s; a[1] /= s; a[2] /= s; }
/* utils.py:48-49: sum([(c2-c1)*(c2-c1) ...]) starting from int 0 */
static double squared_dist(const double *p1, const double *p2) {
    double s = 0.0;
    for (int i = 0; i < 3; ++i) { double d = p1[i] - p2[i]; s = s + d * d; }
    return s;
}

/* utils.py:72-91 */
static int in_triangle(const double *p, const double *tri) {
    const double *v1 = tri, *v2 = tri + 3, *v3 = tri + 6;
    double e[3], q[3], c1[3], c2[3], c3[3];
    sub3(v1, v2, e); sub3(p, v2, q); cross3(e, q, c1);
    sub3(v2, v3, e); sub3(p, v3, q); cross3(e, q, c2);
    sub3(v3, v1, e); sub3(p, v1, q); cross3(e, q, c3);
    div3(c1, norm3(c1)); div3(c2, norm3(c2)); div3(c3, norm3(c3));
    double d12 = dot3(c1, c2), d13 = dot3(c1, c3);
    return (d12 > 0.0) && (d13 > 0.0);   /* np.sign(x) > 0; NaN -> False */
}

/* utils.py:98-147.  `vr` must already be vr/||vr|| (utils.py:110; done once per ray by callers,
 * the value is identical for every triangle).  Returns 1 and P on a hit. */
static int intersect_n(const double *pr, const double *vr, const double *tri, double *P) {
    const double *v1 = tri, *v2 = tri + 3, *v3 = tri + 6;
    double a[3], b[3], n[3];
    sub3(v1, v2, a); sub3(v3, v2, b); cross3(a, b, n);
    div3(n, norm3(n));
    double d = dot3(vr, n);
    if (!(fabs(d) > ZERO)) return 0;
    double t = (dot3(n, v1) - dot3(n, pr)) / dot3(n, vr);
    P[0] = pr[0] + vr[0] * t; P[1] = pr[1] + vr[1] * t; P[2] = pr[2] + vr[2] * t;
    return in_triangle(P, tri);
}

/* single ray-triangle test with the reference's per-call normalisation; for known-answer tests */
int pto_intersect_one(const double *pr, const double *vr_in, const double *tri, double *P) {
    double vr[3] = {vr_in[0], vr_in[1], vr_in[2]};
    div3(vr, norm3(vr));
    return intersect_n(pr, vr, tri, P);
}

/* main.py:83-122: closest hit over objects + light; key = squared distance, first minimum.
 * Returns global triangle index or -1. */
static int closest_hit(const pto_scene *sc, const double *pr, const double *vr_in, double *P_out, double *d2_out) {
    double vr[3] = {vr_in[0], vr_in[1], vr_in[2]};
    div3(vr, norm3(vr));
    int best = -1; double best_d2 = 0.0;
    for (int i = 0; i < sc->n_tri; ++i) {
        double P[3];
        if (!intersect_n(pr, vr, sc->verts + 9 * (size_t)i, P)) continue;
        double d2 = squared_dist(P, pr);
        if (!(d2 > ZERO)) continue;
        if (best < 0 || d2 < best_d2) { best = i; best_d2 = d2; P_out[0] = P[0]; P_out[1] = P[1]; P_out[2] = P[2]; }
    }
    *d2_out = best_d2;
    return best;
}

/* numpy.linspace(start, stop, num)[i]: i*step + start, last element forced to stop */
static double linspace_at(double start, double stop, int num, int i) {
    if (num == 1) return start;
    if (i == num - 1) return stop;
    double step = (stop - start) / (double)(num - 1);
    return (double)i * step + start;
}

/* utils.py:55-69: ray index r = ix*H + iy; origin = eye; direction = (x, y, 0) - eye, unnormalised */
static void primary_ray(const pto_scene *sc, int64_t r, double *o, double *d) {
    int ix = (int)(r / sc->height), iy = (int)(r % sc->height);
    double pt[3] = {linspace_at(sc->ortho[0], sc->ortho[2], sc->width, ix),
                    linspace_at(sc->ortho[1], sc->ortho[3], sc->height, iy), 0.0};
    for (int i = 0; i < 3; ++i) { o[i] = sc->eye[i]; d[i] = pt[i] - sc->eye[i]; }
}

/* parity check 1: primary closest-hit triangle ids */
void pto_primary(const pto_scene *sc, const int64_t *pixels, int64_t n, int32_t *tri, double *d2, double *point) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t i = 0; i < n; ++i) {
        double o[3], d[3], P[3] = {0, 0, 0}, dd = 0;
        primary_ray(sc, pixels ? pixels[i] : i, o, d);
        int t = closest_hit(sc, o, d, P, &dd);
        tri[i] = t;
        if (d2) d2[i] = dd;
        if (point) { point[3 * i] = P[0]; point[3 * i + 1] = P[1]; point[3 * i + 2] = P[2]; }
    }
}

/* generic closest hit for arbitrary rays (known-answer tests) */
void pto_closest(const pto_scene *sc, const double *rays /* [n][6] o,d */, int64_t n, int32_t *tri, double *d2, double *point) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t i = 0; i < n; ++i) {
        double P[3] = {0, 0, 0}, dd = 0;
        tri[i] = closest_hit(sc, rays + 6 * i, rays + 6 * i + 3, P, &dd);
        if (d2) d2[i] = dd;
        if (point) { point[3 * i] = P[0]; point[3 * i + 1] = P[1]; point[3 * i + 2] = P[2]; }
    }
}

/* utils.py:28-39 */
static int pick_random_triangle(const double *areas, int n, double u) {
    double total = 0.0;
    for (int i = 0; i < n; ++i) total = total + areas[i];      /* sum(areas) */
    double x = uniform_ab(0.0, total, u);
    double acc = 0.0;
    for (int i = 0; i < n; ++i) {
        double lo = acc;
        acc = acc + areas[i];
        if (lo <= x && x < acc) return i;
    }
    return -1;   /* reference returns None */
}

/* one light sample (main.py:30-34 -> utils.py:21-46) from the injected stream's dims 4s .. 4s+3: the picked light
 * triangle and the point on it.  Shared by compute_color and the known-answer entry pto_light_sample. */
static int light_sample(const pto_scene *sc, uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, int s, double *lp) {
    const double *lverts = sc->verts + 9 * (size_t)sc->n_tri_obj;
    int li = pick_random_triangle(sc->light_areas, sc->n_light_tri, u01(seed, pixel, sample, bounce, 4 * s));
    double c0 = uniform_ab(0, 1, u01(seed, pixel, sample, bounce, 4 * s + 1));
    double c1 = uniform_ab(0, 1, u01(seed, pixel, sample, bounce, 4 * s + 2));
    double c2 = uniform_ab(0, 1, u01(seed, pixel, sample, bounce, 4 * s + 3));
    double sum = ((0.0 + c0) + c1) + c2;
    double a = c0 / sum, b = c1 / sum, c = c2 / sum;
    const double *lt = lverts + 9 * (size_t)(li < 0 ? 0 : li);
    for (int i = 0; i < 3; ++i) lp[i] = (a * lt[i] + b * lt[3 + i]) + c * lt[6 + i];
    return li;
}

/* main.py:142-145 -> main.py:76-80 + main.py:23-73.  Returns colour; occ/leak are diagnostics. */
static void compute_color(const pto_scene *sc, int obj_index, const double *point, const double *normal,
                          uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce,
                          double *color, int *occ_out, int *leak_out) {
    const double *m = sc->mat + 8 * (size_t)obj_index;
    double amb[3] = {m[0] * m[3] * sc->ambient, m[1] * m[3] * sc->ambient, m[2] * m[3] * sc->ambient};
    double dot = 0.0; int occ = 0; int leak = sc->n_obj - 1;
    for (int s = 0; s < 3; ++s) {
        double lp[3], v[3];
        light_sample(sc, seed, pixel, sample, bounce, s, lp);
        sub3(lp, point, v);
        div3(v, norm3(v));
        double light_d2 = squared_dist(point, lp);
        /* intersect() renormalises the already-normalised direction (utils.py:110) */
        double vr[3] = {v[0], v[1], v[2]};
        div3(vr, norm3(vr));
        int done = 0; leak = sc->n_obj - 1;
        for (int i = 0; i < sc->n_tri_obj; ++i) {
            double P[3];
            if (!intersect_n(point, vr, sc->verts + 9 * (size_t)i, P)) continue;
            double d2 = squared_dist(P, point);
            if (d2 < ZERO) continue;
            if (d2 < light_d2) { done = 1; leak = sc->tri_obj[i]; break; }
        }
        if (done) occ |= 1 << s;
        else dot = dot + dot3(v, normal);
    }
    dot = dot / 3.0;
    const double *lm = sc->mat + 8 * (size_t)leak;   /* leaked `obj` (main.py:70) */
    for (int i = 0; i < 3; ++i) color[i] = amb[i] + sc->light_color[i] * lm[i] * dot;
    *occ_out = occ; *leak_out = leak;
}

/* main.py:148-162 with axis = (0,1,0): the general matrix, its structural zeros kept as zeros */
static void rotate_y(double angle, const double *v, double *out) {
    double a = cos(angle / 2.0);
    double s = sin(angle / 2.0);
    double b = -0.0 * s, c = -1.0 * s, d = -0.0 * s;
    double aa = a * a, bb = b * b, cc = c * c, dd = d * d;
    double bc = b * c, ad = a * d, ac = a * c, ab = a * b, bd = b * d, cd = c * d;
    double R[3][3] = {{aa + bb - cc - dd, 2 * (bc + ad), 2 * (bd - ac)},
                      {2 * (bc - ad), aa + cc - bb - dd, 2 * (cd + ab)},
                      {2 * (bd + ac), 2 * (cd - ab), aa + dd - bb - cc}};
    for (int i = 0; i < 3; ++i) out[i] = dot3(R[i], v);   /* np.dot(matrix, v): same fma chain (probed) */
}

/* one (pixel, sample) path: main.py:186-271 for one value of rays_counter and one pixel.  Outputs indexed
 * [pixel][sample][bounce] like ref_driver.trace; any output pointer may be NULL.  color = the sample's radiance. */
static void trace_sample(const pto_scene *sc, int64_t ip, int64_t r, int s, int32_t spp, int32_t bounces, uint32_t seed,
                         int32_t *tri_out, double *point_out, double *contrib_out, double *k_out,
                         int8_t *lobe_out, int8_t *occ_out, int8_t *leak_out, double *color,
                         int64_t *n_closest, int64_t *n_shadow) {
    double o[3], d[3];
    primary_ray(sc, r, o, d);
    double k = 1.0;
    color[0] = color[1] = color[2] = 0.0;
    int alive = 1;
    for (int b = 0; b < bounces; ++b) {
        size_t idx = ((size_t)ip * spp + s) * bounces + b;
        if (tri_out) tri_out[idx] = -2;
        if (lobe_out) lobe_out[idx] = -1;
        if (occ_out) occ_out[idx] = 0;
        if (leak_out) leak_out[idx] = -1;
        if (k_out) k_out[idx] = 0.0;
        if (point_out) memset(point_out + 3 * idx, 0, 24);
        if (contrib_out) memset(contrib_out + 3 * idx, 0, 24);
        if (!alive) continue;
        double P[3], d2;
        int t = closest_hit(sc, o, d, P, &d2);
        (*n_closest)++;
        if (tri_out) tri_out[idx] = t;
        if (t < 0) { alive = 0; continue; }
        if (point_out) memcpy(point_out + 3 * idx, P, 24);
        int oi = sc->tri_obj[t];
        const double *nrm = sc->normals + 3 * (size_t)t;
        double nc[3];
        if (oi == sc->n_obj) {                      /* light hit: main.py:214-215, 265-266 */
            for (int i = 0; i < 3; ++i) nc[i] = sc->light_color[i];
            alive = 0;
        } else {
            int occ, leak;
            compute_color(sc, oi, P, nrm, seed, (uint32_t)r, (uint32_t)s, (uint32_t)b, nc, &occ, &leak);
            *n_shadow += 3;
            if (occ_out) occ_out[idx] = (int8_t)occ;
            if (leak_out) leak_out[idx] = (int8_t)leak;
        }
        for (int i = 0; i < 3; ++i) {
            double cb = nc[i] * k;
            color[i] = color[i] + cb;               /* main.py:230-231 */
            if (contrib_out) contrib_out[3 * idx + i] = cb;
        }
        if (!alive) { if (k_out) k_out[idx] = k; continue; }
        /* ---- new ray: main.py:236-268 ---- */
        const double *m = sc->mat + 8 * (size_t)oi;
        double kd = m[4], ks = m[5], nexp = m[7];
        double rnd = uniform_ab(0.0, kd + ks, u01(seed, (uint32_t)r, (uint32_t)s, (uint32_t)b, 12));
        double v[3], w[3];
        double angle = acos(nrm[1]);                /* arccos(dot((0,1,0), normal)) */
        if (rnd <= kd) {
            double phi = acos(sqrt(uniform_ab(0, 1, u01(seed, (uint32_t)r, (uint32_t)s, (uint32_t)b, 13))));
            double theta = TAU * uniform_ab(0, 1, u01(seed, (uint32_t)r, (uint32_t)s, (uint32_t)b, 14));
            v[0] = sin(phi) * cos(theta); v[1] = sin(phi) * sin(theta); v[2] = cos(phi);
            div3(v, norm3(v));
            rotate_y(angle, v, w);
            k *= kd * dot3(w, nrm);
            if (lobe_out) lobe_out[idx] = 0;
        } else {
            double s2 = dot3(nrm, d) * 2;
            for (int i = 0; i < 3; ++i) v[i] = s2 * nrm[i] - d[i];
            div3(v, norm3(v));
            double e[3]; sub3(sc->eye, P, e); div3(e, norm3(e));
            rotate_y(angle, v, w);
            k *= ks * pow(dot3(e, w), nexp);
            if (lobe_out) lobe_out[idx] = 1;
        }
        for (int i = 0; i < 3; ++i) { o[i] = P[i]; d[i] = w[i]; }
        if (k_out) k_out[idx] = k;
    }
}

/* full path trace of selected pixels; outputs indexed [pixel][sample][bounce] like ref_driver.trace.
 * Any output pointer except radiance may be NULL.  Threads share the work per pixel when there are many pixels and per
 * (pixel, sample) when there are few (million-triangle scenes are checked on a few dozen pixels); either way a
 * pixel's samples are summed in sample order (main.py:269-271), so the result does not depend on the thread count. */
void pto_trace(const pto_scene *sc, const int64_t *pixels, int64_t n, int32_t spp, int32_t bounces, uint32_t seed,
               int32_t *tri_out, double *point_out, double *contrib_out, double *k_out,
               int8_t *lobe_out, int8_t *occ_out, int8_t *leak_out, double *radiance, int64_t *ray_count) {
    int64_t n_closest = 0, n_shadow = 0;
    int threads = 1;
#ifdef _OPENMP
    threads = omp_get_max_threads();
#endif
    if (n >= 64 * (int64_t)threads) {
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : n_closest, n_shadow)
        for (int64_t ip = 0; ip < n; ++ip) {
            int64_t r = pixels ? pixels[ip] : ip;
            double total[3] = {0, 0, 0};
            for (int s = 0; s < spp; ++s) {
                double color[3];
                trace_sample(sc, ip, r, s, spp, bounces, seed, tri_out, point_out, contrib_out, k_out, lobe_out, occ_out,
                             leak_out, color, &n_closest, &n_shadow);
                for (int i = 0; i < 3; ++i) total[i] = total[i] + color[i];   /* main.py:269-271 */
            }
            for (int i = 0; i < 3; ++i) radiance[3 * ip + i] = total[i] / (double)spp;   /* main.py:277 */
        }
    } else {
        double *colors = (double *)malloc(sizeof(double) * 3 * (size_t)n * (size_t)spp);
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : n_closest, n_shadow)
        for (int64_t q = 0; q < n * (int64_t)spp; ++q) {
            int64_t ip = q / spp;
            int s = (int)(q - ip * spp);
            int64_t r = pixels ? pixels[ip] : ip;
            trace_sample(sc, ip, r, s, spp, bounces, seed, tri_out, point_out, contrib_out, k_out, lobe_out, occ_out,
                         leak_out, colors + 3 * q, &n_closest, &n_shadow);
        }
        for (int64_t ip = 0; ip < n; ++ip) {
            double total[3] = {0, 0, 0};
            for (int s = 0; s < spp; ++s)
                for (int i = 0; i < 3; ++i) total[i] = total[i] + colors[3 * (ip * spp + s) + i];
            for (int i = 0; i < 3; ++i) radiance[3 * ip + i] = total[i] / (double)spp;
        }
        free(colors);
    }
    if (ray_count) { ray_count[0] = n_closest; ray_count[1] = n_shadow; }
}

/* utils.py:150-161.  radiance is [W*H][3] in ray-index order; out is [H][W][3] u8.
 * Index mapping exactly as written (i = counter // width, j = counter % width); like the
 * reference it is only meaningful for width == height. Returns -1 if an index falls outside. */
int pto_make_image(const double *radiance, int32_t width, int32_t height, uint8_t *out) {
    size_t n = (size_t)width * height;
    double *mat = (double *)calloc(n * 3, sizeof(double));
    for (size_t counter = 0; counter < n; ++counter) {
        long i = (long)(counter / width), j = (long)(counter % width);
        long row = height - 1 - j;
        if (row < 0) row += height;
        if (row < 0 || row >= height || i >= width) { free(mat); return -1; }
        for (int c = 0; c < 3; ++c) mat[((size_t)row * width + i) * 3 + c] = radiance[counter * 3 + c];
    }
    double mn = mat[0];
    for (size_t i = 1; i < 3 * n; ++i) if (mat[i] < mn) mn = mat[i];
    for (size_t i = 0; i < 3 * n; ++i) mat[i] = mat[i] - mn;
    double mx = mat[0];
    for (size_t i = 1; i < 3 * n; ++i) if (mat[i] > mx) mx = mat[i];
    for (size_t i = 0; i < 3 * n; ++i) { double v = mat[i] / mx; v = v * 255.0; out[i] = (uint8_t)v; }
    free(mat);
    return 0;
}

/* ---------------- known-answer entries (tests/golden/kat_misc.npz, SURVEY.md 4 tier 1) ---------------- */
/* rotate(axis (0,1,0), angle, v), main.py:148-162 */
void pto_rotate_y(const double *angle, const double *v, int64_t n, double *out) {
    for (int64_t i = 0; i < n; ++i) rotate_y(angle[i], v + 3 * i, out + 3 * i);
}
/* pick_random_triangle + sample_random_pt (utils.py:21-46) at stream position (pixel, sample, bounce, light sample s) */
void pto_light_sample(const pto_scene *sc, uint32_t seed, const int32_t *pos /* [n][4] */, int64_t n, int32_t *pick, double *pt) {
    for (int64_t i = 0; i < n; ++i)
        pick[i] = light_sample(sc, seed, (uint32_t)pos[4 * i], (uint32_t)pos[4 * i + 1], (uint32_t)pos[4 * i + 2], pos[4 * i + 3], pt + 3 * i);
}
/* make_screen_pts + make_rays (utils.py:55-69): direction of ray r as returned (not normalised) */
void pto_primary_dirs(const pto_scene *sc, const int64_t *rays, int64_t n, double *dirs) {
    for (int64_t i = 0; i < n; ++i) { double o[3]; primary_ray(sc, rays[i], o, dirs + 3 * i); }
}

/* bench.py's CPU arm: torchrun exports OMP_NUM_THREADS=1 to its children, which would time one thread and call it
 * "the host"; n <= 0 leaves the runtime's choice alone.  Returns the thread count now in force. */
int pto_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
    return omp_get_max_threads();
#else
    (void)n;
    return 1;
#endif
}

int pto_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
