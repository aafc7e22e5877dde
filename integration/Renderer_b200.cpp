// integration/Renderer_b200.cpp — the binding of INTEGRATION.md §2 as a compilable file.
//
// Drop-in replacement for the reference's src/Renderer.cpp: it defines png::Renderer's two members
// (src/Renderer.h:9-17) against the reference's OWN headers and routes Render() through the C ABI of
// librtm_b200.so (include/rtm_b200.h).  The reference's main.cpp, SettingData.cpp, Ray.h, Renderer.h,
// nlohmann/json and stb stay untouched.  oracle/Makefile compiles it together with the reference's
// unmodified main.cpp + SettingData.cpp (where they lie under /root/reference) into
// oracle/_ref/ref_{head,abc}_b200_cli; tests/test_reference_binding.py runs those next to the
// reference's own CLI.
//
// RTM_COMPAT=head in the environment selects the as-committed semantics, whose image
// is deterministic and must equal the unmodified reference's result.bmp byte for byte.
#include "Renderer.h"   // the reference's

#include <cstdlib>
#include <cstring>
#include <iostream>

#define STB_IMAGE_WRITE_IMPLEMENTATION
#include <stb_image_write.h>   // the reference's vendored copy; its Renderer.cpp is where the implementation lived

#include "rtm_b200.h"

namespace png {

Renderer::Renderer(SettingData& data) : image(std::vector<double>(data.width * data.height * 3)), data(data) {}

void Renderer::Render(std::string fileName) {
    std::vector<double> pos, col, emi;
    std::vector<float> rad;
    for (Object* o : data.object) {   // src/SettingData.h:50; the parser builds spheres only (src/SettingData.cpp:163-179)
        const SphereObject* s = static_cast<const SphereObject*>(o);
        pos.insert(pos.end(), {s->m_position.x, s->m_position.y, s->m_position.z});
        rad.push_back(s->m_size);
        col.insert(col.end(), {s->m_material.color.x, s->m_material.color.y, s->m_material.color.z});
        emi.insert(emi.end(), {s->m_material.emission.x, s->m_material.emission.y, s->m_material.emission.z});
    }
    rtm_scene_desc sc{};
    sc.width = data.width; sc.height = data.height;
    sc.samples = data.samples; sc.superSamples = data.superSamples;
    const vec3 &o = data.camera.origin, &t = data.camera.target, &u = data.camera.upVec;
    const double co[3] = {o.x, o.y, o.z}, ct[3] = {t.x, t.y, t.z}, cu[3] = {u.x, u.y, u.z};
    std::memcpy(sc.cam_origin, co, sizeof co);
    std::memcpy(sc.cam_target, ct, sizeof ct);
    std::memcpy(sc.cam_up, cu, sizeof cu);
    sc.fov = data.camera.fov;
    sc.n_spheres = (int32_t)rad.size();
    sc.position = pos.data(); sc.radius = rad.data(); sc.color = col.data(); sc.emission = emi.data();

    rtm_options opt;
    renderer_default_options(&opt);   // COMPAT_ABC, FP32 bounces, one GPU
    const char* compat = std::getenv("RTM_COMPAT");
    const bool head = compat && std::strcmp(compat, "head") == 0;
    if (head) opt.compat = RTM_COMPAT_HEAD;

    rtm_renderer* r = nullptr;
    if (renderer_create(&sc, &opt, &r) != RTM_OK) {
        std::cerr << "renderer_create: " << renderer_last_error() << std::endl;
        std::exit(1);
    }
    const size_t n = (size_t)data.width * data.height * 3;
    std::vector<float> rgb(n);
    std::vector<unsigned char> resultImage(n);
    int rc = renderer_render(r, rgb.data(), nullptr, nullptr);
    if (rc == RTM_OK && head) rc = renderer_read_image_f64(r, image.data());   // the exact doubles of src/Renderer.cpp:246-248
    if (rc == RTM_OK && !head)
        for (size_t i = 0; i < n; ++i) image[i] = rgb[i];
    if (rc == RTM_OK) rc = renderer_read_u8(r, resultImage.data());             // src/Renderer.cpp:251-254, on the device
    if (rc != RTM_OK) {
        std::cerr << "render: " << renderer_last_error() << std::endl;
        std::exit(1);
    }
    renderer_destroy(r);
    stbi_write_jpg((fileName + ".jpg").c_str(), data.width, data.height, 3, resultImage.data(), 60);   // src/Renderer.cpp:256
    stbi_write_bmp((fileName + ".bmp").c_str(), data.width, data.height, 3, resultImage.data());       // :257
}

}  // namespace png
