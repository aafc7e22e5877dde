/* oracle/rtm_oracle.h — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of RaytracingMin's per-pixel radiance loop, used as
 * the parity checker for the CUDA path.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it; the product
 * (raytracingmin_b200/) never does.
 *
 * Pinning: this restatement is checked bit-for-bit (hit maps AND double images,
 * same per-row mt19937 seeds) against oracle/_ref/libref_{head,abc}.so, which
 * are the reference's own object files (tests/test_oracle_vs_reference.py), and
 * against the golden fixtures under tests/golden/ generated from them.
 */
#ifndef RTM_ORACLE_H
#define RTM_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { RTMO_LEVEL_HEAD = 0, RTMO_LEVEL_ABC = 1 };

/* What SettingData (src/SettingData.h:43-51) holds after parsing, flattened. */
typedef struct rtmo_scene {
    int32_t width, height, samples, superSamples;
    double cam_origin[3], cam_target[3], cam_up[3];
    float fov;               /* src/SettingData.h:45: float, tangent-space half extent */
    int32_t n;               /* number of spheres */
    const double* pos;       /* [n][3]  SphereObject::m_position (double) */
    const float* radius;     /* [n]     SphereObject::m_size     (FLOAT, src/SettingData.h:31) */
    const double* color;     /* [n][3]  Material::color */
    const double* emission;  /* [n][3]  Material::emission */
    /* EXTENSION (no reference semantics to pin): PlaneObject, objectType 2.
     * Object indices n .. n+n_planes-1; ctor arguments of src/SettingData.cpp:235. */
    int32_t n_planes;
    const double* pl_position;  /* [m][3] */
    const double* pl_up;        /* [m][3] */
    const double* pl_target;    /* [m][3] */
    const double* pl_width;     /* [m]    */
    const double* pl_color;     /* [m][3] */
    const double* pl_emission;  /* [m][3] */
} rtmo_scene;

/* Derived plane frame: normal[3], right[3], upp[3], rr, uu (11 doubles), see rtm_oracle.c. */
void rtmo_plane_frame(const double* position3, const double* up3, const double* target3, double width, double* out11);

/* The extension's finite-square test; returns 1 on hit. */
int rtmo_intersect_plane(const double* position3, const double* up3, const double* target3, double width,
                         const double* org3, const double* dir3, double* t_out, double* normal3);

/* src/SettingData.cpp:197-226; returns 1 on hit. normal3 written only on hit. */
int rtmo_intersect_sphere(const double* center3, float radius, const double* org3, const double* dir3,
                          double* t_out, double* normal3);

/* src/Renderer.cpp:202-208: out = camX[3], camY[3], camZ[3], fovx, fovy (11 doubles). */
void rtmo_camera_frame(const rtmo_scene* s, double* out11);

/* src/Renderer.cpp:227-232 */
void rtmo_camera_dir(const rtmo_scene* s, const double* frame11, int x, int y, int sx, int sy, double* dir3);

/* src/Renderer.cpp:58-73: index of the closest hit or -1; t_out/normal3 optional. */
int rtmo_closest_hit(const rtmo_scene* s, const double* org3, const double* dir3, double* t_out, double* normal3);

/* int32[H][W][ss][ss] (axes: y, x, sx-1, sy-1), -1 = miss. */
int rtmo_primary_hits(const rtmo_scene* s, int32_t* out);

/* Linear double image [H][W][3] as src/Renderer.cpp:223-248 accumulates it; the
 * RNG of row y is mt19937(seed + y) read through libstdc++'s
 * uniform_real_distribution<double>(0,1) recipe.  level selects L-HEAD / L-ABC. */
int rtmo_render(const rtmo_scene* s, int level, uint32_t seed, double* image,
                uint64_t* rays_out, uint64_t* samples_out, int nthreads);

/* EXTENSION: the L-ABC estimator with next-event estimation.  The reference's own
 * attempt (SurfaeSample, src/Renderer.cpp:119-198) is dead, unfinished code, so this is a definition, not a
 * restatement; its EXPECTATION equals rtmo_render(level = ABC)'s, which is what the tests check.  Lights =
 * the first RTMO_MAX_LIGHTS spheres in index order with a positive emission component and radius <= 1000.
 * rays_out counts path rays only (comparable with rtmo_render), shadow_out the shadow rays. */
#define RTMO_MAX_LIGHTS 16
int rtmo_render_nee(const rtmo_scene* s, uint32_t seed, double* image, uint64_t* rays_out, uint64_t* samples_out,
                    uint64_t* shadow_out);

/* src/Renderer.cpp:251-254 */
void rtmo_to_u8(const double* image, unsigned char* out, size_t n);

/* Deterministic L-HEAD image (every sample = emission of the primary hit,
 * SURVEY.md F3): no RNG involved, bit-exact target for the GPU's HEAD mode. */
int rtmo_render_head_exact(const rtmo_scene* s, double* image);

#ifdef __cplusplus
}
#endif
#endif
