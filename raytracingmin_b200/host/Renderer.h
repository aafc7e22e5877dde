// Renderer.h — the host façade with the reference's interface (src/Renderer.h:9-17):
// construct from a SettingData reference, call Render(fileName) -> fileName.jpg + fileName.bmp.
// Render() is the binding of INTEGRATION.md §2: flatten -> renderer_create ->
// renderer_render -> renderer_read_u8 -> renderer_destroy -> image writers.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "SettingData.h"

namespace png {

struct RenderOptions {           // everything the reference does not have; defaults = drop-in behaviour
    Compat compat = Compat::Abc;
    int device = 0, n_gpus = 1;
    int precision = 0;           // RTM_PRECISION_F32
    int max_depth = 0;
    int accel = 0;               // RTM_ACCEL_AUTO: brute force up to 64 spheres, the hierarchy beyond
    int variant = 0;             // RTM_VARIANT_MEGA (1 = wavefront, kept for comparison)
    int estimator = 0;           // RTM_ESTIMATOR_REFERENCE (1 = next-event estimation, extension)
    uint64_t seed = 0x5EED;
    bool want_hits = false;      // keep the primary-hit map (int32[H][W][ss][ss])
    bool quiet = false;
};

struct RenderStats {
    uint64_t samples = 0, rays = 0, sphere_tests = 0, bvh_nodes = 0;
    double device_ms = 0, e2e_ms = 0;
    int kernel_launches = 0, n_gpus = 0;
};

class Renderer {
public:
    explicit Renderer(SettingData& data) : Renderer(data, RenderOptions{}) {}
    Renderer(SettingData& data, const RenderOptions& opt) : options(opt), data(data) {
        image.resize((size_t)data.width * data.height * 3);
    }
    void Render(std::string fileName);

    RenderOptions options;
    RenderStats stats;
    std::vector<double> image;        // [H][W][3] linear sums, as src/Renderer.h:15
    std::vector<unsigned char> image8;  // [H][W][3] after src/Renderer.cpp:251-254
    std::vector<int32_t> hits;        // filled when options.want_hits
private:
    SettingData& data;
};

}  // namespace png
