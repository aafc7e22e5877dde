#include "Renderer.h"

#include <cstdio>
#include <iostream>
#include <stdexcept>

#include "../../include/rtm_b200.h"
#include "image_write.hpp"

namespace png {

void Renderer::Render(std::string fileName) {
    rtm_scene_desc sc{};
    sc.width = data.width; sc.height = data.height;
    sc.samples = data.samples; sc.superSamples = data.superSamples;
    const double o[3] = {data.camera.origin.x, data.camera.origin.y, data.camera.origin.z};
    const double t[3] = {data.camera.target.x, data.camera.target.y, data.camera.target.z};
    const double u[3] = {data.camera.upVec.x, data.camera.upVec.y, data.camera.upVec.z};
    for (int k = 0; k < 3; ++k) { sc.cam_origin[k] = o[k]; sc.cam_target[k] = t[k]; sc.cam_up[k] = u[k]; }
    sc.fov = data.camera.fov;
    sc.n_spheres = (int32_t)data.object.size();
    sc.position = data.object.position.data();
    sc.radius = data.object.radius.data();
    sc.color = data.object.color.data();
    sc.emission = data.object.emission.data();
    sc.n_planes = (int32_t)data.plane.size();   // PlaneObject extension (ABI 3)
    if (sc.n_planes > 0) {
        sc.plane_position = data.plane.position.data(); sc.plane_up = data.plane.up.data();
        sc.plane_target = data.plane.target.data(); sc.plane_width = data.plane.width.data();
        sc.plane_color = data.plane.color.data(); sc.plane_emission = data.plane.emission.data();
    }

    rtm_options opt;
    renderer_default_options(&opt);
    opt.device = options.device; opt.n_gpus = options.n_gpus;
    opt.compat = (options.compat == Compat::Head) ? RTM_COMPAT_HEAD : RTM_COMPAT_ABC;
    opt.precision = options.precision; opt.max_depth = options.max_depth; opt.seed = options.seed;
    opt.accel = options.accel; opt.variant = options.variant; opt.estimator = options.estimator;

    rtm_renderer* r = nullptr;
    if (renderer_create(&sc, &opt, &r) != RTM_OK) throw std::runtime_error(std::string("renderer_create: ") + renderer_last_error());
    const size_t n_pix = (size_t)data.width * data.height;
    std::vector<float> rgb(n_pix * 3);
    if (options.want_hits) hits.resize(n_pix * data.superSamples * data.superSamples);
    rtm_stats st{};
    if (!options.quiet) std::cout << "0 / " << data.height << std::endl;   // progress, cf. src/Renderer.cpp:216
    int rc = renderer_render(r, rgb.data(), options.want_hits ? hits.data() : nullptr, &st);
    if (rc == RTM_OK) {
        image8.resize(n_pix * 3);
        rc = renderer_read_u8(r, image8.data());
    }
    if (rc == RTM_OK && options.compat == Compat::Head) rc = renderer_read_image_f64(r, image.data());  // exact doubles
    const std::string err = (rc == RTM_OK) ? "" : renderer_last_error();
    renderer_destroy(r);
    if (rc != RTM_OK) throw std::runtime_error("render: " + err);
    if (options.compat != Compat::Head)
        for (size_t i = 0; i < rgb.size(); ++i) image[i] = rgb[i];
    if (!options.quiet) std::cout << data.height << " / " << data.height << std::endl;
    stats.samples = st.samples; stats.rays = st.rays; stats.device_ms = st.device_ms; stats.e2e_ms = st.e2e_ms;
    stats.kernel_launches = st.kernel_launches; stats.n_gpus = st.n_gpus;
    stats.sphere_tests = st.sphere_tests; stats.bvh_nodes = st.bvh_nodes;

    // src/Renderer.cpp:256-257: JPEG quality 60 and BMP, same names
    if (!imgw::write_jpg(fileName + ".jpg", image8.data(), data.width, data.height, 60))
        throw std::runtime_error("cannot write " + fileName + ".jpg");
    if (!imgw::write_bmp(fileName + ".bmp", image8.data(), data.width, data.height))
        throw std::runtime_error("cannot write " + fileName + ".bmp");
}

}  // namespace png
