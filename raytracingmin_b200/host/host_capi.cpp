// host_capi.cpp — a tiny C surface over the host-side pieces (scene loader, image
// writers) so that the CPU test-suite can exercise them through ctypes without a GPU.
#include <cstring>
#include <string>

#include "SettingData.h"
#include "image_write.hpp"

namespace { thread_local std::string g_err; }

extern "C" {

const char* rtmh_last_error(void) { return g_err.c_str(); }

// Parses a scene file.  Scalars -> dims[4] = {width,height,samples,superSamples}, cam[9], fov.
// Returns the sphere count (>= 0) or -1.  Arrays may be NULL to query the count first.
int rtmh_load_scene(const char* path, int compat_head, int* dims, double* cam, float* fov, int cap,
                    double* position, float* radius, double* color, double* emission) {
    try {
        png::LoadData ld(path, compat_head ? png::Compat::Head : png::Compat::Abc);
        const png::SettingData& d = ld.data;
        if (dims) { dims[0] = d.width; dims[1] = d.height; dims[2] = d.samples; dims[3] = d.superSamples; }
        if (cam) {
            const double c[9] = {d.camera.origin.x, d.camera.origin.y, d.camera.origin.z, d.camera.target.x, d.camera.target.y,
                                 d.camera.target.z, d.camera.upVec.x, d.camera.upVec.y, d.camera.upVec.z};
            std::memcpy(cam, c, sizeof c);
        }
        if (fov) *fov = d.camera.fov;
        const int n = (int)d.object.size();
        if (position && n > 0 && n <= cap) {   // n == 0: empty vectors hand memcpy a null source (UB even for 0 bytes)
            std::memcpy(position, d.object.position.data(), sizeof(double) * 3 * n);
            std::memcpy(radius, d.object.radius.data(), sizeof(float) * n);
            std::memcpy(color, d.object.color.data(), sizeof(double) * 3 * n);
            std::memcpy(emission, d.object.emission.data(), sizeof(double) * 3 * n);
        }
        return n;
    } catch (const std::exception& e) {
        g_err = e.what();
        return -1;
    }
}

// The planes of a scene file (extension, objectType 2).  Returns the plane count or -1; arrays may be NULL.
// order_out[n_spheres + n_planes]: object id -> position in the JSON list.
int rtmh_load_planes(const char* path, int cap, double* position, double* up, double* target, double* width, double* color,
                     double* emission, int order_cap, int* order_out) {
    try {
        png::LoadData ld(path, png::Compat::Abc);
        const png::PlaneList& P = ld.data.plane;
        const int m = (int)P.size();
        if (position && m > 0 && m <= cap) {
            std::memcpy(position, P.position.data(), sizeof(double) * 3 * m);
            std::memcpy(up, P.up.data(), sizeof(double) * 3 * m);
            std::memcpy(target, P.target.data(), sizeof(double) * 3 * m);
            std::memcpy(width, P.width.data(), sizeof(double) * m);
            std::memcpy(color, P.color.data(), sizeof(double) * 3 * m);
            std::memcpy(emission, P.emission.data(), sizeof(double) * 3 * m);
        }
        if (order_out && !ld.data.order.empty() && (int)ld.data.order.size() <= order_cap)
            std::memcpy(order_out, ld.data.order.data(), sizeof(int) * ld.data.order.size());
        return m;
    } catch (const std::exception& e) {
        g_err = e.what();
        return -1;
    }
}

// Round trip: parse `path`, serialise with to_json_text into out (NUL-terminated); returns length or -1.
int rtmh_roundtrip_json(const char* path, char* out, int cap) {
    try {
        png::LoadData ld(path, png::Compat::Abc);
        const std::string s = png::LoadData::to_json_text(ld.data);
        if ((int)s.size() + 1 > cap) { g_err = "buffer too small"; return -1; }
        std::memcpy(out, s.c_str(), s.size() + 1);
        return (int)s.size();
    } catch (const std::exception& e) {
        g_err = e.what();
        return -1;
    }
}

int rtmh_save_sample_json(const char* path) { png::LoadData::SaveSampleJson(path); return 0; }
int rtmh_write_bmp(const char* path, const unsigned char* rgb, int w, int h) { return imgw::write_bmp(path, rgb, w, h) ? 0 : -1; }
int rtmh_write_jpg(const char* path, const unsigned char* rgb, int w, int h, int quality) {
    return imgw::write_jpg(path, rgb, w, h, quality) ? 0 : -1;
}

}  // extern "C"
