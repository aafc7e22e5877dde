// host_sanitize_driver.cpp — TEST INFRASTRUCTURE.  Drives the host front end's C surface (raytracingmin_b200/host/host_capi.cpp:
// scene loader with the superset policy, JSON round trip, sample-JSON writer, BMP/JPEG writers) inside ONE executable built
// with -fsanitize=address,undefined (scripts/host_asan.sh).  A Python process with a preloaded ASan runtime aborts on the first
// C++ exception thrown inside a dlopen'ed library, so the sanitised run goes through this driver instead of pytest.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

extern "C" {
const char* rtmh_last_error(void);
int rtmh_load_scene(const char* path, int compat_head, int* dims, double* cam, float* fov, int cap, double* position, float* radius,
                    double* color, double* emission);
int rtmh_load_planes(const char* path, int cap, double* position, double* up, double* target, double* width, double* color,
                     double* emission, int order_cap, int* order_out);
int rtmh_roundtrip_json(const char* path, char* out, int cap);
int rtmh_save_sample_json(const char* path);
int rtmh_write_bmp(const char* path, const unsigned char* rgb, int w, int h);
int rtmh_write_jpg(const char* path, const unsigned char* rgb, int w, int h, int quality);
}

static int fails = 0;
#define EXPECT(c)                                                          \
    do {                                                                   \
        if (!(c)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #c); ++fails; } \
    } while (0)

static int load(const std::string& path, bool head, std::vector<double>& pos) {
    int dims[4]; double cam[9]; float fov;
    const int n = rtmh_load_scene(path.c_str(), head, dims, cam, &fov, 0, nullptr, nullptr, nullptr, nullptr);
    if (n < 0) return n;
    pos.assign(3 * (size_t)n + 3, 0.0);
    std::vector<float> rad(n + 1);
    std::vector<double> col(3 * (size_t)n + 3), emi(3 * (size_t)n + 3);
    return rtmh_load_scene(path.c_str(), head, dims, cam, &fov, n, pos.data(), rad.data(), col.data(), emi.data());
}

int main(int argc, char** argv) {
    const std::string scenes = argc > 1 ? argv[1] : "scenes", tmp = argc > 2 ? argv[2] : "/tmp";
    std::vector<double> pos;
    for (const char* nm : {"settingData.head.json", "cornellBox.head.json", "simple1.head.json", "simple2.head.json", "synth10k.json",
                           "cornellBoxPlanes.json"})
        for (int head = 0; head < 2; ++head) {
            const int n = load(scenes + "/" + nm, head, pos);
            const bool planes_in_head = head && std::strstr(nm, "Planes");      // planes are an extension of L-ABC: a clean error
            if (planes_in_head) { EXPECT(n < 0 && std::strlen(rtmh_last_error()) > 0); continue; }
            if (n < 3) std::printf("  %s head=%d -> %d (%s)\n", nm, head, n, rtmh_last_error());
            EXPECT(n >= 3);
            std::vector<char> buf(8 << 20);
            EXPECT(rtmh_roundtrip_json((scenes + "/" + nm).c_str(), buf.data(), (int)buf.size()) > 0);
        }
    {   // planes
        double p[16 * 3], u[16 * 3], t[16 * 3], w[16], c[16 * 3], e[16 * 3];
        int order[64];
        EXPECT(rtmh_load_planes((scenes + "/cornellBoxPlanes.json").c_str(), 16, p, u, t, w, c, e, 64, order) == 6);
    }
    // malformed input: every one must come back as an error code with a message, never as a crash
    const char* bad[] = {"", "{", "{\"00 width\": 8,", "[1,2,3]", "{\"00 width\": \"x\"}", "{\"00 width\":8,\"00 height\":4,\"00 samples\":1}",
                         "{\"00 width\":8,\"00 height\":4,\"00 samples\":1,\"01 camera\":{\"origin\":[0,0],\"target\":[0,0,0],\"upVec\":[0,1,0],\"fov\":1}}",
                         "{\"00 width\":8,\"00 height\":4,\"00 samples\":1,\"01 camera\":{\"origin\":[0,0,-1],\"target\":[0,0,0],\"upVec\":[0,1,0],\"fov\":1},"
                         "\"02 scene\":{\"00 object\":[{\"00 objectType\":3,\"00 position\":[0,0,0],\"01 size\":1,\"02 material\":{\"color\":[1,1,1],\"emission\":[0,0,0]}}]}}",
                         "{\"a\": \"\\u12\"}", "{\"a\": 1e99999}", "\xff\xfe\x00"};
    for (const char* b : bad) {
        const std::string f = tmp + "/rtm_bad.json";
        std::ofstream(f, std::ios::binary) << b;
        const int n = load(f, false, pos);
        EXPECT(n < 0 && std::strlen(rtmh_last_error()) > 0);
    }
    EXPECT(load(tmp + "/does_not_exist.json", false, pos) < 0);
    EXPECT(rtmh_save_sample_json((tmp + "/rtm_sample.json").c_str()) == 0);
    EXPECT(load(tmp + "/rtm_sample.json", false, pos) >= 0);
    // writers on awkward sizes (1x1, not multiples of 8 / 16, one row)
    for (auto wh : {std::pair<int, int>{1, 1}, {7, 3}, {17, 33}, {64, 48}, {250, 1}}) {
        std::vector<unsigned char> img((size_t)wh.first * wh.second * 3);
        for (size_t i = 0; i < img.size(); ++i) img[i] = (unsigned char)((i * 2654435761u) >> 13);
        EXPECT(rtmh_write_bmp((tmp + "/rtm_t.bmp").c_str(), img.data(), wh.first, wh.second) == 0);
        for (int q : {1, 60, 90, 100}) EXPECT(rtmh_write_jpg((tmp + "/rtm_t.jpg").c_str(), img.data(), wh.first, wh.second, q) == 0);
    }
    std::printf("host sanitize driver: %d failure(s)\n", fails);
    return fails ? 1 : 0;
}
