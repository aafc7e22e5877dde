// SettingData.h — the host front end's scene/config types.  They play the role of the
// reference's png::SettingData / png::Camera / png::LoadData (src/SettingData.h:43-61)
// with the same member names for the scalar settings, so that code written against the
// reference (`loadData.data.width`, `data.camera.fov`, ...) keeps compiling; the object
// list is flat arrays ready for the C ABI instead of a vector of polymorphic Object*.
#pragma once
#include <string>
#include <vector>

namespace png {

struct vec3 {
    double x = 0, y = 0, z = 0;
    vec3() = default;
    vec3(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
};

struct Camera {
    vec3 origin, target, upVec;
    float fov = 0.f;  // FLOAT as in src/SettingData.h:45; tangent-space half width (src/Renderer.cpp:207,229)
};

// Spheres only: the reference's parser builds nothing else (src/SettingData.cpp:163-179).
struct SphereList {
    std::vector<double> position;  // [n][3]  SphereObject::m_position
    std::vector<float> radius;     // [n]     SphereObject::m_size (float, src/SettingData.h:31)
    std::vector<double> color;     // [n][3]  Material::color
    std::vector<double> emission;  // [n][3]  Material::emission
    size_t size() const { return radius.size(); }
};

// EXTENSION: the reference's PlaneObject ("00 objectType": 2, src/SettingData.h:33-42), which its
// parser never builds and its Intersect never finishes.  Flat arrays of the ctor arguments of
// src/SettingData.cpp:235 (position, up, target, width) + material.  Planes take the object ids
// behind the spheres; `order` gives every id's position in the JSON object list.
struct PlaneList {
    std::vector<double> position, up, target;  // [m][3]
    std::vector<double> width;                 // [m]
    std::vector<double> color, emission;       // [m][3]
    size_t size() const { return width.size(); }
};

struct SettingData {
    int width = 0, height = 0, samples = 0, superSamples = 0;
    Camera camera;
    SphereList object;
    PlaneList plane;
    std::vector<int> order;  // object id (spheres first, then planes) -> index in "00 object" ({} entries not counted)
};

enum class Compat { Head, Abc };

struct SceneError : std::exception {
    std::string msg;
    explicit SceneError(std::string m) : msg(std::move(m)) {}
    const char* what() const noexcept override { return msg.c_str(); }
};

class LoadData {
public:
    // Parses `jsonFile` (src/SettingData.cpp:6-12).  Loader policy = superset of the
    // reference's from_json (src/SettingData.cpp:129-186), SURVEY.md §5.1:
    //   unknown keys ignored; "00 sample" accepted for "00 samples"; missing
    //   "00 superSamples" = 1; object without "00 objectType" = sphere; `{}` skipped;
    //   objectType 2 = plane (extension: keys "00 position", "01 up", "01 target", "01 width",
    //   "02 material"; Compat::Abc only); any other objectType rejected with a message (the
    //   reference stores a null object and segfaults);
    //   Compat::Head applies the reference's position bug (posi.x assigned three times,
    //   src/SettingData.cpp:165-167): every sphere lands at (position[2], 0, 0).
    explicit LoadData(const std::string& jsonFile, Compat compat = Compat::Abc);
    SettingData data;

    static SettingData from_json_text(const std::string& text, Compat compat);
    // Full serialisation in the HEAD schema (the reference's to_json leaves the scene
    // out: its loop body is commented out, src/SettingData.cpp:119-127).
    static std::string to_json_text(const SettingData& data);
    // The bootstrap file of src/SettingData.cpp:14-104: 960x540, 10 samples, ss 4, camera
    // at the origin looking down +z, fov 60, no objects.
    static void SaveSampleJson(const std::string& fileName);
};

}  // namespace png
