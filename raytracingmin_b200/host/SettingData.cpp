#include "SettingData.h"

#include <fstream>
#include <sstream>

#include "json_min.hpp"

namespace png {
namespace {

using jsonmin::Value;

vec3 read_vec3(const Value& v, const std::string& what) {
    if (!v.is_array() || v.size() != 3) throw SceneError(what + ": expected an array of 3 numbers");
    return vec3(v.at(0).number(), v.at(1).number(), v.at(2).number());
}

int read_int(const Value& v, const std::string& what) {
    if (!v.is_number()) throw SceneError(what + ": number expected");
    return (int)v.number();
}

const Value& need(const Value& obj, const std::string& key, const std::string& where) {
    if (!obj.is_object() || !obj.contains(key)) throw SceneError(where + ": missing key \"" + key + "\"");
    return obj.at(key);
}

Value vec3_json(double x, double y, double z) {
    Value a = Value::array();
    a.push_back(Value(x)); a.push_back(Value(y)); a.push_back(Value(z));
    return a;
}

}  // namespace

SettingData LoadData::from_json_text(const std::string& text, Compat compat) {
    Value root;
    try {
        root = jsonmin::parse(text);
    } catch (const std::exception& e) {
        throw SceneError(e.what());
    }
    if (!root.is_object()) throw SceneError("scene file: top level must be an object");
    SettingData d;
    try {
        d.width = read_int(need(root, "00 width", "scene"), "00 width");
        d.height = read_int(need(root, "00 height", "scene"), "00 height");
        if (root.contains("00 samples")) d.samples = read_int(root.at("00 samples"), "00 samples");
        else if (root.contains("00 sample")) d.samples = read_int(root.at("00 sample"), "00 sample");  // legacy key
        else throw SceneError("scene: missing key \"00 samples\"");
        d.superSamples = root.contains("00 superSamples") ? read_int(root.at("00 superSamples"), "00 superSamples") : 1;
        const Value& cam = need(root, "01 camera", "scene");
        d.camera.origin = read_vec3(need(cam, "origin", "01 camera"), "camera origin");
        d.camera.target = read_vec3(need(cam, "target", "01 camera"), "camera target");
        d.camera.upVec = read_vec3(need(cam, "upVec", "01 camera"), "camera upVec");
        d.camera.fov = (float)need(cam, "fov", "01 camera").number();
        if (root.contains("02 scene") && root.at("02 scene").contains("00 object")) {
            const Value& objs = root.at("02 scene").at("00 object");
            if (!objs.is_array()) throw SceneError("\"00 object\" must be an array");
            std::vector<int> order_s, order_p;
            int k = 0;
            for (size_t i = 0; i < objs.size(); ++i) {
                const Value& o = objs.at(i);
                const std::string where = "object " + std::to_string(i);
                if (o.empty()) continue;  // `{}` (cornellBoxSetting.json:47): skipped, index not consumed
                const int type = o.contains("00 objectType") ? read_int(o.at("00 objectType"), where + " objectType") : 1;
                if (type == 2) {
                    if (compat == Compat::Head) throw SceneError(where + ": planes are an extension of the L-ABC level");
                    const vec3 pp = read_vec3(need(o, "00 position", where), where + " position");
                    const vec3 pu = read_vec3(need(o, "01 up", where), where + " up");
                    const vec3 pt = read_vec3(need(o, "01 target", where), where + " target");
                    const double w = need(o, "01 width", where).number();
                    if (!(w > 0.0)) throw SceneError(where + ": plane width must be positive");
                    const Value& pm = need(o, "02 material", where);
                    const vec3 pc = read_vec3(need(pm, "color", where + " material"), where + " color");
                    const vec3 pe = read_vec3(need(pm, "emission", where + " material"), where + " emission");
                    d.plane.position.insert(d.plane.position.end(), {pp.x, pp.y, pp.z});
                    d.plane.up.insert(d.plane.up.end(), {pu.x, pu.y, pu.z});
                    d.plane.target.insert(d.plane.target.end(), {pt.x, pt.y, pt.z});
                    d.plane.width.push_back(w);
                    d.plane.color.insert(d.plane.color.end(), {pc.x, pc.y, pc.z});
                    d.plane.emission.insert(d.plane.emission.end(), {pe.x, pe.y, pe.z});
                    order_p.push_back(k++);
                    continue;
                }
                if (type != 1)
                    throw SceneError(where + ": objectType " + std::to_string(type) +
                                     " is neither a sphere (1) nor a plane (2) (the reference stores a null object and crashes)");
                order_s.push_back(k++);
                vec3 p = read_vec3(need(o, "00 position", where), where + " position");
                if (compat == Compat::Head) p = vec3(p.z, 0.0, 0.0);  // F2
                const float size = (float)need(o, "01 size", where).number();
                const Value& mat = need(o, "02 material", where);
                const vec3 c = read_vec3(need(mat, "color", where + " material"), where + " color");
                const vec3 e = read_vec3(need(mat, "emission", where + " material"), where + " emission");
                d.object.position.insert(d.object.position.end(), {p.x, p.y, p.z});
                d.object.radius.push_back(size);
                d.object.color.insert(d.object.color.end(), {c.x, c.y, c.z});
                d.object.emission.insert(d.object.emission.end(), {e.x, e.y, e.z});
            }
            d.order = order_s;
            d.order.insert(d.order.end(), order_p.begin(), order_p.end());
        }
    } catch (const SceneError&) {
        throw;
    } catch (const std::exception& e) {
        throw SceneError(e.what());
    }
    if (d.width <= 0 || d.height <= 0 || d.samples <= 0 || d.superSamples <= 0)
        throw SceneError("width, height, samples and superSamples must be positive");
    return d;
}

LoadData::LoadData(const std::string& jsonFile, Compat compat) {
    std::ifstream in(jsonFile, std::ios::binary);
    if (!in) throw SceneError("cannot open " + jsonFile);
    std::stringstream ss;
    ss << in.rdbuf();
    data = from_json_text(ss.str(), compat);
}

std::string LoadData::to_json_text(const SettingData& d) {
    Value root = Value::object();
    root["00 width"] = Value(d.width);
    root["00 height"] = Value(d.height);
    root["00 samples"] = Value(d.samples);
    root["00 superSamples"] = Value(d.superSamples);
    Value cam = Value::object();
    cam["origin"] = vec3_json(d.camera.origin.x, d.camera.origin.y, d.camera.origin.z);
    cam["target"] = vec3_json(d.camera.target.x, d.camera.target.y, d.camera.target.z);
    cam["upVec"] = vec3_json(d.camera.upVec.x, d.camera.upVec.y, d.camera.upVec.z);
    cam["fov"] = Value((double)d.camera.fov);
    root["01 camera"] = cam;
    if (d.object.size() + d.plane.size() > 0) {
        Value objs = Value::array();
        for (size_t i = 0; i < d.object.size(); ++i) {
            Value o = Value::object();
            o["00 objectType"] = Value(1);
            o["00 position"] = vec3_json(d.object.position[3 * i], d.object.position[3 * i + 1], d.object.position[3 * i + 2]);
            o["01 size"] = Value((double)d.object.radius[i]);
            Value m = Value::object();
            m["color"] = vec3_json(d.object.color[3 * i], d.object.color[3 * i + 1], d.object.color[3 * i + 2]);
            m["emission"] = vec3_json(d.object.emission[3 * i], d.object.emission[3 * i + 1], d.object.emission[3 * i + 2]);
            o["02 material"] = m;
            objs.push_back(o);
        }
        for (size_t j = 0; j < d.plane.size(); ++j) {
            const PlaneList& P = d.plane;
            Value o = Value::object();
            o["00 objectType"] = Value(2);
            o["00 position"] = vec3_json(P.position[3 * j], P.position[3 * j + 1], P.position[3 * j + 2]);
            o["01 up"] = vec3_json(P.up[3 * j], P.up[3 * j + 1], P.up[3 * j + 2]);
            o["01 target"] = vec3_json(P.target[3 * j], P.target[3 * j + 1], P.target[3 * j + 2]);
            o["01 width"] = Value(P.width[j]);
            Value m = Value::object();
            m["color"] = vec3_json(P.color[3 * j], P.color[3 * j + 1], P.color[3 * j + 2]);
            m["emission"] = vec3_json(P.emission[3 * j], P.emission[3 * j + 1], P.emission[3 * j + 2]);
            o["02 material"] = m;
            objs.push_back(o);
        }
        Value scene = Value::object();
        scene["00 object"] = objs;
        root["02 scene"] = scene;
    }
    return jsonmin::dump(root);
}

void LoadData::SaveSampleJson(const std::string& fileName) {
    SettingData d;
    d.width = 960; d.height = 540; d.samples = 10; d.superSamples = 4;
    d.camera.origin = vec3(0, 0, 0);
    d.camera.target = vec3(0, 0, 1);
    d.camera.upVec = vec3(0, 1, 0);
    d.camera.fov = 60.f;
    std::ofstream out(fileName, std::ios::binary);
    out << to_json_text(d);
}

}  // namespace png
