/* oracle/rtm_oracle.c — TEST INFRASTRUCTURE ONLY (see rtm_oracle.h).
 *
 * Plain-C restatement of the reference's radiance loop.  Every function cites
 * the reference lines it follows (paths relative to /root/reference).  Compile
 * with -ffp-contract=off and without -march=native so that no FMA is formed:
 * the primary-hit index is compared bit-for-bit.
 *
 * The float quirks that matter are kept on purpose:
 *   - Magnitude() rounds the squared length to float and takes a FLOAT sqrt,
 *   - the sphere radius is float and r*r is a float product,
 *   - minValue is the float 1e-5f, `rate` is float, kd() returns float.
 */
#include "rtm_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { double x, y, z; } v3;

static v3 v3_make(double x, double y, double z) { v3 r = {x, y, z}; return r; }
static v3 v3_add(v3 a, v3 b) { return v3_make(a.x + b.x, a.y + b.y, a.z + b.z); }
static v3 v3_sub(v3 a, v3 b) { return v3_make(a.x - b.x, a.y - b.y, a.z - b.z); }
static v3 v3_neg(v3 a) { return v3_make(-a.x, -a.y, -a.z); }
static v3 v3_mul(v3 a, v3 b) { return v3_make(a.x * b.x, a.y * b.y, a.z * b.z); }
static v3 v3_scale(v3 a, double s) { return v3_make(a.x * s, a.y * s, a.z * s); }
static v3 v3_div(v3 a, double s) { return v3_make(a.x / s, a.y / s, a.z / s); }

/* src/Ray.h:61-63 */
static double dot3(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* src/Ray.h:64-66 (note the y component is written -ax*bz + az*bx) */
static v3 cross3(v3 a, v3 b) { return v3_make(a.y * b.z - a.z * b.y, -a.x * b.z + a.z * b.x, a.x * b.y - a.y * b.x); }
/* src/Ray.h:67-69: sqrtf on the double sum => float argument, float sqrt */
static double magnitude3(v3 a) { return (double)sqrtf((float)(a.x * a.x + a.y * a.y + a.z * a.z)); }
/* src/Ray.h:70-72: three divides */
static v3 normalize3(v3 a) { return v3_div(a, magnitude3(a)); }

static v3 v3_load(const double* p) { return v3_make(p[0], p[1], p[2]); }
static void v3_store(double* p, v3 a) { p[0] = a.x; p[1] = a.y; p[2] = a.z; }

/* ---- sphere (src/SettingData.cpp:197-226) -------------------------------- */
static int intersect_sphere(v3 center, float radius, v3 org, v3 dir, double* out_dis, v3* out_normal) {
    const v3 p_o = v3_sub(center, org);
    const double b = dot3(p_o, dir);
    const float r2 = radius * radius; /* float x float, rounded to float (:200) */
    const double D4 = b * b - dot3(p_o, p_o) + r2;
    if (D4 < 0.0) return 0;
    const double sqrt_D4 = sqrt(D4);
    const double t1 = b - sqrt_D4, t2 = b + sqrt_D4;
    const float minValue = 1e-5f;
    if (t1 < minValue && t2 < minValue) return 0;
    *out_dis = (t1 > 0.001) ? t1 : t2;
    {
        const v3 hitPoint = v3_add(org, v3_scale(dir, *out_dis));
        *out_normal = normalize3(v3_sub(hitPoint, center));
    }
    return 1;
}

int rtmo_intersect_sphere(const double* center3, float radius, const double* org3, const double* dir3,
                          double* t_out, double* normal3) {
    double t = 0.0;
    v3 n = {0, 0, 0};
    const int hit = intersect_sphere(v3_load(center3), radius, v3_load(org3), v3_load(dir3), &t, &n);
    *t_out = t;
    if (hit && normal3) v3_store(normal3, n);
    return hit;
}

/* ---- plane: EXTENSION -----------------------------------------
 * The reference declares PlaneObject (src/SettingData.h:33-42) but its Intersect only
 * rejects parallel rays and then falls off the end (src/SettingData.cpp:243-246): there is
 * nothing to be in parity with.  Defined here as a finite square, keeping what the
 * reference does define:
 *   ctor (src/SettingData.cpp:235-242): normal = Normalize(target - position) (float-sqrt
 *        Normalize), right = Normalize(Cross(normal, up)) * 0.5 * width;
 *   added: upp = Cross(right, normal) (half-extent vector of the other side),
 *        rr = Dot(right,right), uu = Dot(upp,upp);
 *   Intersect: reject |Dot(normal,dir)| < FLT_EPSILON (:244-245); t = Dot(normal, position
 *        - org) / Dot(normal, dir); accept t > 0.001 (the sphere's near threshold, :209) with
 *        q = org + dir*t - position, |Dot(q,right)| <= rr and |Dot(q,upp)| <= uu;
 *        out_normal = normal (PathTracing orients it, src/Renderer.cpp:82-83). */
typedef struct { v3 position, normal, right, upp; double rr, uu; } plane_t;

static plane_t make_plane(v3 position, v3 up, v3 target, double width) {
    plane_t p;
    p.position = position;
    p.normal = normalize3(v3_sub(target, position));
    p.right = v3_scale(v3_scale(normalize3(cross3(p.normal, up)), 0.5), width);
    p.upp = cross3(p.right, p.normal);
    p.rr = dot3(p.right, p.right);
    p.uu = dot3(p.upp, p.upp);
    return p;
}

static int intersect_plane(const plane_t* p, v3 org, v3 dir, double* out_dis, v3* out_normal) {
    const double dn = dot3(p->normal, dir);
    if (fabs(dn) < FLT_EPSILON) return 0;
    const double t = dot3(p->normal, v3_sub(p->position, org)) / dn;
    if (!(t > 0.001)) return 0;
    const v3 q = v3_sub(v3_add(org, v3_scale(dir, t)), p->position);
    if (fabs(dot3(q, p->right)) > p->rr || fabs(dot3(q, p->upp)) > p->uu) return 0;
    *out_dis = t;
    *out_normal = p->normal;
    return 1;
}

static plane_t scene_plane(const rtmo_scene* s, int j) {
    return make_plane(v3_load(s->pl_position + 3 * j), v3_load(s->pl_up + 3 * j), v3_load(s->pl_target + 3 * j), s->pl_width[j]);
}

void rtmo_plane_frame(const double* position3, const double* up3, const double* target3, double width, double* out11) {
    const plane_t p = make_plane(v3_load(position3), v3_load(up3), v3_load(target3), width);
    v3_store(out11, p.normal); v3_store(out11 + 3, p.right); v3_store(out11 + 6, p.upp);
    out11[9] = p.rr; out11[10] = p.uu;
}

int rtmo_intersect_plane(const double* position3, const double* up3, const double* target3, double width,
                         const double* org3, const double* dir3, double* t_out, double* normal3) {
    const plane_t p = make_plane(v3_load(position3), v3_load(up3), v3_load(target3), width);
    double t = 0.0;
    v3 n = {0, 0, 0};
    const int hit = intersect_plane(&p, v3_load(org3), v3_load(dir3), &t, &n);
    *t_out = t;
    if (hit && normal3) v3_store(normal3, n);
    return hit;
}

/* ---- camera (src/Renderer.cpp:202-208, 227-232) --------------------------- */
typedef struct { v3 camX, camY, camZ; double fovx, fovy; } frame_t;

static frame_t make_frame(const rtmo_scene* s) {
    frame_t f;
    const v3 direction = normalize3(v3_sub(v3_load(s->cam_target), v3_load(s->cam_origin)));
    f.camX = v3_neg(normalize3(cross3(direction, v3_load(s->cam_up))));
    f.camY = cross3(f.camX, direction);
    f.camZ = direction;
    f.fovx = (double)s->fov;
    f.fovy = f.fovx * s->height / s->width;
    return f;
}

static v3 camera_dir(const rtmo_scene* s, const frame_t* f, int x, int y, int sx, int sy) {
    const float rate = (float)(1.0 / (1 + s->superSamples));
    /* rate*sx is float*int -> float; 2.0f*(double) -> double; /int -> double */
    const double px = 2.0f * ((double)x + rate * sx) / s->width - 1.0f;
    const double py = 2.0f * ((double)y + rate * sy) / s->height - 1.0f;
    const v3 a = v3_scale(v3_scale(f->camX, f->fovx), px);
    const v3 b = v3_scale(v3_scale(f->camY, f->fovy), py);
    return normalize3(v3_add(v3_add(a, b), f->camZ));
}

void rtmo_camera_frame(const rtmo_scene* s, double* out11) {
    const frame_t f = make_frame(s);
    v3_store(out11, f.camX); v3_store(out11 + 3, f.camY); v3_store(out11 + 6, f.camZ);
    out11[9] = f.fovx; out11[10] = f.fovy;
}

void rtmo_camera_dir(const rtmo_scene* s, const double* fr, int x, int y, int sx, int sy, double* dir3) {
    frame_t f;
    f.camX = v3_load(fr); f.camY = v3_load(fr + 3); f.camZ = v3_load(fr + 6);
    f.fovx = fr[9]; f.fovy = fr[10];
    v3_store(dir3, camera_dir(s, &f, x, y, sx, sy));
}

/* ---- closest hit (src/Renderer.cpp:58-73) --------------------------------- */
static int closest_hit(const rtmo_scene* s, v3 org, v3 dir, double* dis_out, v3* normal_out) {
    int hitObject = -1;
    double dis = DBL_MAX;
    v3 normal = {0, 0, 0};
    for (int i = 0; i < s->n; ++i) {
        double tmp_dis;
        v3 tmp_normal;
        if (intersect_sphere(v3_load(s->pos + 3 * i), s->radius[i], org, dir, &tmp_dis, &tmp_normal) &&
            tmp_dis < dis && tmp_dis > 0) { /* strict <: lowest index wins ties */
            dis = tmp_dis;
            normal = tmp_normal;
            hitObject = i;
        }
    }
    for (int j = 0; j < s->n_planes; ++j) { /* extension: planes follow the spheres in the object list */
        double tmp_dis;
        v3 tmp_normal;
        const plane_t p = scene_plane(s, j);
        if (intersect_plane(&p, org, dir, &tmp_dis, &tmp_normal) && tmp_dis < dis && tmp_dis > 0) {
            dis = tmp_dis;
            normal = tmp_normal;
            hitObject = s->n + j;
        }
    }
    if (dis_out) *dis_out = dis;
    if (normal_out) *normal_out = normal;
    return hitObject;
}

int rtmo_closest_hit(const rtmo_scene* s, const double* org3, const double* dir3, double* t_out, double* normal3) {
    double t;
    v3 n;
    const int h = closest_hit(s, v3_load(org3), v3_load(dir3), &t, &n);
    if (t_out) *t_out = t;
    if (normal3) v3_store(normal3, n);
    return h;
}

int rtmo_primary_hits(const rtmo_scene* s, int32_t* out) {
    const frame_t f = make_frame(s);
    const int ss = s->superSamples;
    const v3 org = v3_load(s->cam_origin);
#pragma omp parallel for schedule(dynamic)
    for (int y = 0; y < s->height; ++y)
        for (int x = 0; x < s->width; ++x)
            for (int sx = 1; sx <= ss; ++sx)
                for (int sy = 1; sy <= ss; ++sy)
                    out[(((size_t)y * s->width + x) * ss + (sx - 1)) * ss + (sy - 1)] =
                        closest_hit(s, org, camera_dir(s, &f, x, y, sx, sy), NULL, NULL);
    return 0;
}

/* ---- RNG: std::mt19937 + uniform_real_distribution<double>(0,1) ------------
 * The reference draws from one mt19937 through a uniform_real_distribution
 * (src/Renderer.cpp:210-213).  Restated: the standard MT19937 recurrence, and
 * libstdc++'s generate_canonical<double,53> (two 32-bit outputs per double:
 * (a + b*2^32) / 2^64, clamped below 1).  Per-row seeding matches
 * oracle/ref_harness.cpp so images can be compared bit-for-bit. */
typedef struct { uint32_t mt[624]; int idx; } mt19937_t;

static void mt_seed(mt19937_t* g, uint32_t seed) {
    g->mt[0] = seed;
    for (int i = 1; i < 624; ++i) g->mt[i] = 1812433253u * (g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) + (uint32_t)i;
    g->idx = 624;
}
static uint32_t mt_next(mt19937_t* g) {
    if (g->idx >= 624) {
        for (int i = 0; i < 624; ++i) {
            const uint32_t y = (g->mt[i] & 0x80000000u) | (g->mt[(i + 1) % 624] & 0x7fffffffu);
            g->mt[i] = g->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        g->idx = 0;
    }
    uint32_t y = g->mt[g->idx++];
    y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
    return y;
}
static double mt_uniform01(mt19937_t* g) {
    const double a = (double)mt_next(g);
    const double b = (double)mt_next(g);
    double r = (a + b * 4294967296.0) / 18446744073709551616.0;
    if (r >= 1.0) r = nextafter(1.0, 0.0);
    return r;
}

/* ---- material (src/SettingData.h:8-17) ------------------------------------ */
static float material_kd(v3 color) { /* returns FLOAT */
    const double m = (color.x < color.y) ? color.y : color.x; /* std::max(a,b) = (a<b)?b:a */
    return (float)((m < color.z) ? color.z : m);
}

/* ---- PathTracing (src/Renderer.cpp:57-117) -------------------------------- */
typedef struct { uint64_t rays; } counters_t;

static v3 path_trace(const rtmo_scene* s, int level, v3 org, v3 dir, mt19937_t* rng, counters_t* c) {
    const double PI = 3.14159265358979323846; /* std::numbers::pi */
    double dis;
    v3 normal;
    c->rays++;
    const int hitObject = closest_hit(s, org, dir, &dis, &normal);
    if (hitObject == -1 || !(dis > 0)) return v3_make(0, 0, 0);
    const int isPlane = hitObject >= s->n;
    const v3 color = isPlane ? v3_load(s->pl_color + 3 * (hitObject - s->n)) : v3_load(s->color + 3 * hitObject);
    const v3 emission = isPlane ? v3_load(s->pl_emission + 3 * (hitObject - s->n)) : v3_load(s->emission + 3 * hitObject);
    const float kd = material_kd(color);
    if (!(mt_uniform01(rng) <= kd)) return emission; /* :78, :112 */
    const v3 colorKD = v3_div(color, (double)kd);      /* SettingData.h:11-13 */
    if (level == RTMO_LEVEL_HEAD) {
        /* F3: the normal never reaches the caller => bounce basis is NaN => the
         * recursive call misses everything and returns (0,0,0) before any draw;
         * the call still counts as a ray (it runs the closest-hit loop), and the two
         * draws of :88-89 are still consumed from the stream.  :108-109 */
        (void)mt_uniform01(rng);
        (void)mt_uniform01(rng);
        c->rays++;
        return v3_add(v3_mul(colorKD, v3_make(0, 0, 0)), emission);
    }
    const v3 hitPoint = v3_add(v3_scale(dir, dis), org);                       /* :79 */
    const v3 w = (dot3(normal, dir) < 0.0) ? normal : v3_scale(normal, -1.0); /* :82-83,87 */
    const double r1 = 2 * PI * mt_uniform01(rng);                             /* :88 */
    const double r2 = mt_uniform01(rng);                                      /* :89 */
    const double r2s = sqrt(r2);                                              /* :90 */
    v3 u;
    if (fabs(w.x) > FLT_MIN) u = normalize3(cross3(v3_make(0, 1, 0), w));     /* :96-101 */
    else u = normalize3(cross3(v3_make(1, 0, 0), w));
    const v3 v = cross3(w, u);                                                /* :102 */
    const v3 nextDir = normalize3(v3_add(v3_add(v3_scale(v3_scale(u, cos(r1)), r2s),
                                                v3_scale(v3_scale(v, sin(r1)), r2s)),
                                         v3_scale(w, sqrt(1.0 - r2))));       /* :103-107 */
    const v3 L = path_trace(s, level, hitPoint, nextDir, rng, c);             /* :108 (F4 fixed in L-ABC) */
    return v3_add(v3_mul(colorKD, L), emission);                              /* :109 */
}

/* src/Renderer.cpp:43-49 with std::min/std::max NaN behaviour */
static double clamp01(double v) {
    const double lo = (v < 0.0) ? 0.0 : v;          /* std::max<double>(v, 0) */
    const double hi = (double)1.0f;
    return (hi < lo) ? hi : lo;                     /* std::min<double>(lo, 1) */
}

int rtmo_render(const rtmo_scene* s, int level, uint32_t seed, double* image,
                uint64_t* rays_out, uint64_t* samples_out, int nthreads) {
    const frame_t f = make_frame(s);
    const v3 org = v3_load(s->cam_origin);
    uint64_t rays = 0, samples = 0;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();   /* an explicit count overrides OMP_NUM_THREADS (torchrun sets it to 1) */
#else
    (void)nthreads;
#endif
#pragma omp parallel for schedule(dynamic) num_threads(nthreads) reduction(+ : rays, samples)
    for (int y = 0; y < s->height; ++y) {
        mt19937_t rng;
        counters_t c = {0};
        mt_seed(&rng, seed + (uint32_t)y);
        for (int x = 0; x < s->width; ++x) {
            v3 acc = {0, 0, 0};
            for (int sx = 1; sx <= s->superSamples; ++sx)
                for (int sy = 1; sy <= s->superSamples; ++sy)
                    for (int smp = 0; smp < s->samples; ++smp) {
                        const v3 dir = camera_dir(s, &f, x, y, sx, sy);
                        v3 cal = {0, 0, 0};
                        if (mt_uniform01(&rng) >= 1.0) { /* :234 strategy pick: never taken */
                        } else {
                            cal = path_trace(s, level, org, dir, &rng, &c);
                        }
                        cal = v3_div(v3_div(v3_div(cal, s->superSamples), s->superSamples), s->samples); /* :240 */
                        cal = v3_make(clamp01(cal.x), clamp01(cal.y), clamp01(cal.z));                   /* :241 */
                        acc = v3_add(acc, cal);
                        ++samples;
                    }
            v3_store(image + ((size_t)y * s->width + x) * 3, acc);
        }
        rays += c.rays;
    }
    if (rays_out) *rays_out = rays;
    if (samples_out) *samples_out = samples;
    return 0;
}

/* ---- next-event estimation: EXTENSION ---------------------------------------
 * Same path as path_trace (L-ABC): closest hit, roulette on kd, cosine-weighted bounce.  Added at every
 * vertex that survives the roulette: one light k chosen uniformly from the list, a direction l sampled
 * uniformly inside the cone the light sphere subtends (pdf 1 / (2 pi (1 - cos_max))), a shadow ray, and
 *     direct = e_k * max(0, w.l) * 2 (1 - cos_max) * n_lights        (Lambert: colorKD * cos / pi / pdf)
 * when the shadow ray's closest hit is light k.  To count every emitted photon once, a hit reached by a
 * BOUNCE ray adds its emission only if it was not eligible for the light sampling at the previous vertex:
 * not in the list, or hit from inside (the previous vertex was inside the sphere: no cone).  The primary
 * hit always adds its emission.  A vertex ON light k (hitObject == k) skips the sampling: a convex light
 * does not illuminate itself. */
typedef struct { int n; int id[RTMO_MAX_LIGHTS]; unsigned char listed[1 << 16]; } lights_t;

static void make_lights(const rtmo_scene* s, lights_t* L) {
    L->n = 0;
    memset(L->listed, 0, sizeof L->listed);
    for (int i = 0; i < s->n && L->n < RTMO_MAX_LIGHTS; ++i) {
        const double* e = s->emission + 3 * i;
        if ((e[0] > 0 || e[1] > 0 || e[2] > 0) && s->radius[i] <= 1000.0f) {
            L->id[L->n++] = i;
            if (i < (1 << 16)) L->listed[i] = 1;
        }
    }
}

typedef struct { uint64_t rays, shadow; } nee_counters_t;

static v3 path_trace_nee(const rtmo_scene* s, const lights_t* lights, int primary, v3 org, v3 dir, mt19937_t* rng,
                         nee_counters_t* c) {
    const double PI = 3.14159265358979323846;
    double dis;
    v3 normal;
    c->rays++;
    const int hitObject = closest_hit(s, org, dir, &dis, &normal);
    if (hitObject == -1 || !(dis > 0)) return v3_make(0, 0, 0);
    const int isPlane = hitObject >= s->n;
    const v3 color = isPlane ? v3_load(s->pl_color + 3 * (hitObject - s->n)) : v3_load(s->color + 3 * hitObject);
    v3 emission = isPlane ? v3_load(s->pl_emission + 3 * (hitObject - s->n)) : v3_load(s->emission + 3 * hitObject);
    const int from_inside = !(dot3(normal, dir) < 0.0);
    if (!primary && !isPlane && hitObject < (1 << 16) && lights->listed[hitObject] && !from_inside)
        emission = v3_make(0, 0, 0); /* already counted by the light sampling of the previous vertex */
    const float kd = material_kd(color);
    if (!(mt_uniform01(rng) <= kd)) return emission;
    const v3 colorKD = v3_div(color, (double)kd);
    const v3 hitPoint = v3_add(v3_scale(dir, dis), org);
    const v3 w = from_inside ? v3_scale(normal, -1.0) : normal;
    v3 direct = v3_make(0, 0, 0);
    if (lights->n > 0) {
        const double uk = mt_uniform01(rng), u1 = mt_uniform01(rng), u2 = mt_uniform01(rng);
        int k = (int)(uk * lights->n);
        if (k >= lights->n) k = lights->n - 1;
        const int lid = lights->id[k];
        const v3 v = v3_sub(v3_load(s->pos + 3 * lid), hitPoint);
        const double d2 = dot3(v, v), r2 = (double)s->radius[lid] * (double)s->radius[lid];
        if (lid != hitObject && d2 > r2) {
            const double cos_max = sqrt(1.0 - r2 / d2);
            const double omc = (r2 / d2) / (1.0 + cos_max); /* 1 - cos_max without cancellation */
            const double ct = 1.0 - u1 * omc, st = sqrt(u1 * omc * (2.0 - u1 * omc));
            const v3 a = v3_scale(v, 1.0 / sqrt(d2));
            v3 t1;
            if (fabs(a.x) > 0.5) t1 = cross3(v3_make(0, 1, 0), a);
            else t1 = cross3(v3_make(1, 0, 0), a);
            t1 = v3_scale(t1, 1.0 / sqrt(dot3(t1, t1)));
            const v3 t2 = cross3(a, t1);
            const double ph = 2 * PI * u2;
            const v3 l = v3_add(v3_scale(a, ct), v3_scale(v3_add(v3_scale(t1, cos(ph)), v3_scale(t2, sin(ph))), st));
            const double cw = dot3(w, l);
            if (cw > 0.0) {
                double sd;
                v3 sn;
                c->shadow++;
                if (closest_hit(s, hitPoint, l, &sd, &sn) == lid)
                    direct = v3_scale(v3_load(s->emission + 3 * lid), cw * 2.0 * omc * lights->n);
            }
        }
    }
    const double r1 = 2 * PI * mt_uniform01(rng);
    const double r2 = mt_uniform01(rng);
    const double r2s = sqrt(r2);
    v3 u;
    if (fabs(w.x) > FLT_MIN) u = normalize3(cross3(v3_make(0, 1, 0), w));
    else u = normalize3(cross3(v3_make(1, 0, 0), w));
    const v3 vv = cross3(w, u);
    const v3 nextDir = normalize3(v3_add(v3_add(v3_scale(v3_scale(u, cos(r1)), r2s), v3_scale(v3_scale(vv, sin(r1)), r2s)),
                                         v3_scale(w, sqrt(1.0 - r2))));
    const v3 L = path_trace_nee(s, lights, 0, hitPoint, nextDir, rng, c);
    return v3_add(v3_mul(colorKD, v3_add(direct, L)), emission);
}

int rtmo_render_nee(const rtmo_scene* s, uint32_t seed, double* image, uint64_t* rays_out, uint64_t* samples_out,
                    uint64_t* shadow_out) {
    const frame_t f = make_frame(s);
    const v3 org = v3_load(s->cam_origin);
    uint64_t rays = 0, samples = 0, shadow = 0;
    lights_t* lights = (lights_t*)malloc(sizeof(lights_t));
    make_lights(s, lights);
#pragma omp parallel for schedule(dynamic) reduction(+ : rays, samples, shadow)
    for (int y = 0; y < s->height; ++y) {
        mt19937_t rng;
        nee_counters_t c = {0, 0};
        mt_seed(&rng, seed + (uint32_t)y);
        for (int x = 0; x < s->width; ++x) {
            v3 acc = {0, 0, 0};
            for (int sx = 1; sx <= s->superSamples; ++sx)
                for (int sy = 1; sy <= s->superSamples; ++sy)
                    for (int smp = 0; smp < s->samples; ++smp) {
                        (void)mt_uniform01(&rng); /* the strategy-pick draw of src/Renderer.cpp:234, kept so that an
                                                     empty light list reproduces rtmo_render(ABC) bit for bit */
                        v3 cal = path_trace_nee(s, lights, 1, org, camera_dir(s, &f, x, y, sx, sy), &rng, &c);
                        cal = v3_div(v3_div(v3_div(cal, s->superSamples), s->superSamples), s->samples);
                        cal = v3_make(clamp01(cal.x), clamp01(cal.y), clamp01(cal.z));
                        acc = v3_add(acc, cal);
                        ++samples;
                    }
            v3_store(image + ((size_t)y * s->width + x) * 3, acc);
        }
        rays += c.rays;
        shadow += c.shadow;
    }
    free(lights);
    if (rays_out) *rays_out = rays;
    if (samples_out) *samples_out = samples;
    if (shadow_out) *shadow_out = shadow;
    return 0;
}

int rtmo_render_head_exact(const rtmo_scene* s, double* image) {
    const frame_t f = make_frame(s);
    const v3 org = v3_load(s->cam_origin);
#pragma omp parallel for schedule(dynamic)
    for (int y = 0; y < s->height; ++y)
        for (int x = 0; x < s->width; ++x) {
            v3 acc = {0, 0, 0};
            for (int sx = 1; sx <= s->superSamples; ++sx)
                for (int sy = 1; sy <= s->superSamples; ++sy) {
                    const int h = closest_hit(s, org, camera_dir(s, &f, x, y, sx, sy), NULL, NULL);
                    v3 cal = (h < 0) ? v3_make(0, 0, 0)
                                     : (h >= s->n) ? v3_load(s->pl_emission + 3 * (h - s->n)) : v3_load(s->emission + 3 * h);
                    cal = v3_div(v3_div(v3_div(cal, s->superSamples), s->superSamples), s->samples);
                    cal = v3_make(clamp01(cal.x), clamp01(cal.y), clamp01(cal.z));
                    for (int smp = 0; smp < s->samples; ++smp) acc = v3_add(acc, cal);
                }
            v3_store(image + ((size_t)y * s->width + x) * 3, acc);
        }
    return 0;
}

/* src/Renderer.cpp:251-254: (unsigned char)255 * min(v, 1.0) -> int*double,
 * then double -> unsigned char truncation. */
void rtmo_to_u8(const double* image, unsigned char* out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        const double v = (1.0 < image[i]) ? 1.0 : image[i]; /* std::min(image, 1.0) */
        out[i] = (unsigned char)((unsigned char)255 * v);
    }
}
