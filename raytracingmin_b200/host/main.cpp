// main.cpp — command line of the host front end; same flags, defaults and side effects as
// the reference's src/main.cpp:8-46:
//   -?            usage, exit 0
//   -json <file>  scene file; a missing argument or one that starts with '-' selects the default
//   -sampleJson   write the sample settingData.json and exit 0
// default file settingData.json, written first if it does not exist; output result.jpg +
// result.bmp in the working directory.  Everything new carries a "--" prefix so it cannot
// collide with a reference flag.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>
#include <sys/stat.h>

#include "Renderer.h"
#include "SettingData.h"

namespace {

bool file_exists(const std::string& p) {
    struct stat st;
    return ::stat(p.c_str(), &st) == 0;
}

bool opt_value(const std::string& arg, const char* name, std::string& out) {
    const std::string pre = std::string(name) + "=";
    if (arg.compare(0, pre.size(), pre) != 0) return false;
    out = arg.substr(pre.size());
    return true;
}

void usage() {
    std::cout << "Usage: rtm_b200 [OPTION]...\n\n"
              << "-sampleJson : write the sample scene file settingData.json and exit\n"
              << "-json <file> : scene file to render (default settingData.json)\n"
              << "--compat=abc|head  --gpus=N  --device=N  --seed=N  --precision=f32|f64  --max-depth=N\n"
              << "--accel=auto|brute|bvh  --variant=mega|wavefront  --estimator=reference|nee\n"
              << "--width=N --height=N --samples=N --superSamples=N : override the file's settings\n"
              << "--out=<name> : output base name (default result)  --dump-hits=<file>  --stats  --quiet\n"
              << "--dump-pfm=<file> : the linear image before the 8-bit conversion (float RGB, Portable Float Map)\n";
}

}  // namespace

int main(int argc, char* argv[]) {
    std::string jsonFile = "settingData.json", outName = "result", dumpHits, dumpPfm, v;
    png::RenderOptions ro;
    int ow = 0, oh = 0, os = 0, oss = 0;
    bool stats = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "-?") { usage(); return 0; }
        else if (a == "-json") {
            // like the reference, the file name is NOT consumed: it is examined again as a flag and ignored
            if (i + 1 < argc && argv[i + 1][0] != '-') jsonFile = argv[i + 1];
            else jsonFile = "settingData.json";
        } else if (a == "-sampleJson") {
            std::cout << "writing the sample scene file settingData.json" << std::endl;
            png::LoadData::SaveSampleJson("settingData.json");
            return 0;
        } else if (opt_value(a, "--compat", v)) ro.compat = (v == "head") ? png::Compat::Head : png::Compat::Abc;
        else if (opt_value(a, "--gpus", v)) ro.n_gpus = std::atoi(v.c_str());
        else if (opt_value(a, "--device", v)) ro.device = std::atoi(v.c_str());
        else if (opt_value(a, "--seed", v)) ro.seed = std::strtoull(v.c_str(), nullptr, 0);
        else if (opt_value(a, "--precision", v)) ro.precision = (v == "f64") ? 1 : 0;
        else if (opt_value(a, "--max-depth", v)) ro.max_depth = std::atoi(v.c_str());
        else if (opt_value(a, "--accel", v)) ro.accel = (v == "brute") ? 1 : (v == "bvh") ? 2 : 0;
        else if (opt_value(a, "--variant", v)) ro.variant = (v == "wavefront") ? 1 : 0;
        else if (opt_value(a, "--estimator", v)) ro.estimator = (v == "nee") ? 1 : 0;
        else if (opt_value(a, "--width", v)) ow = std::atoi(v.c_str());
        else if (opt_value(a, "--height", v)) oh = std::atoi(v.c_str());
        else if (opt_value(a, "--samples", v)) os = std::atoi(v.c_str());
        else if (opt_value(a, "--superSamples", v)) oss = std::atoi(v.c_str());
        else if (opt_value(a, "--out", v)) outName = v;
        else if (opt_value(a, "--dump-hits", v)) { dumpHits = v; ro.want_hits = true; }
        else if (opt_value(a, "--dump-pfm", v)) dumpPfm = v;
        else if (a == "--stats") stats = true;
        else if (a == "--quiet") ro.quiet = true;
    }
    if (!file_exists(jsonFile)) {  // bootstrap, src/main.cpp:36-39
        std::cout << "writing the sample scene file " << jsonFile << std::endl;
        png::LoadData::SaveSampleJson(jsonFile);
    }
    std::cout << "loading " << jsonFile << " and rendering" << std::endl;
    try {
        png::LoadData loadData(jsonFile, ro.compat);
        if (ow > 0) loadData.data.width = ow;
        if (oh > 0) loadData.data.height = oh;
        if (os > 0) loadData.data.samples = os;
        if (oss > 0) loadData.data.superSamples = oss;
        png::Renderer renderer(loadData.data, ro);
        renderer.Render(outName);
        if (!dumpPfm.empty()) {
            // what the RMSE tooling compares ("linear float images before 8-bit conversion"; the
            // JPEG is lossy, the BMP quantised and clipped at 1.0).  PFM: rows bottom-to-top, negative scale = little endian.
            std::ofstream f(dumpPfm, std::ios::binary);
            f << "PF\n" << loadData.data.width << " " << loadData.data.height << "\n-1.0\n";
            std::vector<float> row((size_t)loadData.data.width * 3);
            for (int y = loadData.data.height - 1; y >= 0; --y) {
                for (size_t i = 0; i < row.size(); ++i) row[i] = (float)renderer.image[(size_t)y * row.size() + i];
                f.write(reinterpret_cast<const char*>(row.data()), (std::streamsize)(row.size() * sizeof(float)));
            }
        }
        if (!dumpHits.empty()) {
            std::ofstream f(dumpHits, std::ios::binary);
            f.write(reinterpret_cast<const char*>(renderer.hits.data()), (std::streamsize)(renderer.hits.size() * sizeof(int32_t)));
        }
        if (stats) {
            const auto& s = renderer.stats;
            std::printf("{\"samples\": %llu, \"rays\": %llu, \"device_ms\": %.3f, \"e2e_ms\": %.3f, \"mrays_per_s\": %.1f, "
                        "\"msamples_per_s\": %.1f, \"kernel_launches\": %d, \"n_gpus\": %d, \"sphere_tests\": %llu, \"bvh_nodes\": %llu}\n",
                        (unsigned long long)s.samples, (unsigned long long)s.rays, s.device_ms, s.e2e_ms,
                        s.rays / (s.device_ms * 1e3), s.samples / (s.device_ms * 1e3), s.kernel_launches, s.n_gpus,
                        (unsigned long long)s.sphere_tests, (unsigned long long)s.bvh_nodes);
        }
    } catch (const std::exception& e) {
        std::cerr << "error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
