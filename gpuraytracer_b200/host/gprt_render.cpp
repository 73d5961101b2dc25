// gprt_render -- the headless stand-in for the reference's main.cpp + Window message loop
// (src/main.cpp:9-20, src/Window.cpp:107-119): build a scene, run on_init once, then per frame
// on_update + on_render with frames in flight like the reference's swap chain (frame_count = 3,
// src/Renderer.h:92), and instead of Present write the frames as binary PPM files.  Prints what the
// reference puts in its window title (calculate_frame_stats, src/Renderer.cpp:375-399): frames per
// second and million primary rays per second, here over the whole run.
//
//   gprt_render [--scene ID] [--width W --height H] [--grid N] [--frames K] [--time T0] [--dt DT]
//               [--camera-orbit DEG] [--light-orbit DEG] [--device D | --devices D0,D1,...] [--in-flight F] [--out PREFIX]
//
// --devices cuts every frame into interleaved row stripes over the listed CUDA devices (one context each, one process;
// a device may be listed more than once) and every device copies its rows straight into the one page-locked host frame
// the PPM is written from (gprt_multi_*, host/multi_renderer.h).
//
// Links against libgprt.so only (no CUDA toolkit needed on the include path).  There is no CPU
// render path: without an sm_100 GPU the program fails with GPRT_E_NO_DEVICE.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/gprt_host.h"

namespace {

struct Options {
    int scene = 0, device = 0, frames = 1, in_flight = 3;
    uint32_t width = 0, height = 0, grid = 0;
    float t0 = 0.0f, dt = 1.0f / 60.0f, camera_orbit = 0.0f, light_orbit = 0.0f;
    std::string out;
    std::vector<int> devices; // --devices: one frame over several devices
};

void usage() {
    std::puts("gprt_render [--scene 0..7] [--width W --height H] [--grid N] [--frames K] [--time T0] [--dt DT]\n"
              "            [--camera-orbit DEG_PER_FRAME] [--light-orbit DEG_PER_FRAME] [--device D | --devices D0,D1,...] [--in-flight 1..8] [--out PREFIX]\n"
              "scenes: 0 reference scene, 1..5 the BASELINE configs, 6 config 3 with the long march budget, 7 instanced TLAS");
}

bool parse(int argc, char** argv, Options& o) {
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&](const char* what) -> const char* {
            if (i + 1 >= argc) { std::fprintf(stderr, "gprt_render: %s needs a value\n", what); return nullptr; }
            return argv[++i];
        };
        const char* v = nullptr;
        if (a == "--help" || a == "-h") { usage(); std::exit(0); }
        else if (a == "--scene") { if (!(v = next("--scene"))) return false; o.scene = std::atoi(v); }
        else if (a == "--width") { if (!(v = next("--width"))) return false; o.width = (uint32_t)std::atoi(v); }
        else if (a == "--height") { if (!(v = next("--height"))) return false; o.height = (uint32_t)std::atoi(v); }
        else if (a == "--grid") { if (!(v = next("--grid"))) return false; o.grid = (uint32_t)std::atoi(v); }
        else if (a == "--frames") { if (!(v = next("--frames"))) return false; o.frames = std::atoi(v); }
        else if (a == "--time") { if (!(v = next("--time"))) return false; o.t0 = (float)std::atof(v); }
        else if (a == "--dt") { if (!(v = next("--dt"))) return false; o.dt = (float)std::atof(v); }
        else if (a == "--camera-orbit") { if (!(v = next("--camera-orbit"))) return false; o.camera_orbit = (float)std::atof(v); }
        else if (a == "--light-orbit") { if (!(v = next("--light-orbit"))) return false; o.light_orbit = (float)std::atof(v); }
        else if (a == "--device") { if (!(v = next("--device"))) return false; o.device = std::atoi(v); }
        else if (a == "--devices") {
            if (!(v = next("--devices"))) return false;
            for (const char* p = v; *p;) { o.devices.push_back(std::atoi(p)); while (*p && *p != ',') p++; if (*p == ',') p++; }
        }
        else if (a == "--in-flight") { if (!(v = next("--in-flight"))) return false; o.in_flight = std::atoi(v); }
        else if (a == "--out") { if (!(v = next("--out"))) return false; o.out = v; }
        else { std::fprintf(stderr, "gprt_render: unknown option %s\n", a.c_str()); return false; }
    }
    return o.frames >= 1 && o.in_flight >= 1 && o.in_flight <= 8;
}

int fail(const char* what, int rc) {
    std::fprintf(stderr, "gprt_render: %s failed: %s (%d)\n", what, gprt_error_string(rc), rc);
    return 1;
}

} // namespace

int main(int argc, char** argv) {
    Options o;
    if (!parse(argc, argv, o)) { usage(); return 2; }
    gprt_renderer* r = nullptr;
    gprt_multi* m = nullptr;
    int rc;
    uint32_t w = 0, h = 0;
    if (!o.devices.empty()) {
        rc = gprt_multi_create(&m, o.scene, o.width, o.height, o.grid, o.devices.data(), (int)o.devices.size(), nullptr);
        if (rc != GPRT_OK) return fail("gprt_multi_create", rc);
        if ((rc = gprt_multi_on_init(m)) != GPRT_OK) { gprt_multi_destroy(m); return fail("on_init", rc); }
        if ((rc = gprt_multi_set_frame_count(m, o.in_flight)) != GPRT_OK) { gprt_multi_destroy(m); return fail("set_frame_count", rc); }
        gprt_multi_get_size(m, &w, &h);
    } else {
        rc = gprt_renderer_create(&r, o.scene, o.width, o.height, o.grid, o.device, nullptr);
        if (rc != GPRT_OK) return fail("gprt_renderer_create", rc);
        if ((rc = gprt_renderer_on_init(r)) != GPRT_OK) { gprt_renderer_destroy(r); return fail("on_init", rc); }
        if ((rc = gprt_renderer_set_frame_count(r, o.in_flight)) != GPRT_OK) { gprt_renderer_destroy(r); return fail("set_frame_count", rc); }
        gprt_renderer_get_size(r, &w, &h);
    }
    const size_t frame_bytes = (size_t)w * h * 4;
    std::vector<void*> host(o.in_flight, nullptr);
    std::vector<int> frame_of(o.in_flight, -1);
    for (auto& p : host)
        if (!(p = gprt_alloc_host(frame_bytes))) { std::fprintf(stderr, "gprt_render: no pinned memory for a %ux%u frame\n", w, h); return 1; }

    auto write_frame = [&](int slot) -> int {
        if (o.out.empty() || frame_of[slot] < 0) return GPRT_OK;
        char name[512];
        std::snprintf(name, sizeof name, "%s%04d.ppm", o.out.c_str(), frame_of[slot]);
        return gprt_write_ppm(name, host[slot], w, h);
    };

    // Frame statistics over the steady state: the first frames of every slot allocate that slot's queues and page in the
    // kernels, so the clock starts (after a full wait) once each slot has rendered twice -- if the run is long enough.
    const int warm = (o.frames >= 4 * o.in_flight) ? 2 * o.in_flight : 0;
    auto t_begin = std::chrono::steady_clock::now();
    for (int f = 0; f < o.frames && rc == GPRT_OK; f++) {
        if (f == warm && warm) {
            rc = m ? gprt_multi_wait_for_gpu(m) : gprt_renderer_wait_for_gpu(r);
            if (rc != GPRT_OK) break;
            t_begin = std::chrono::steady_clock::now();
        }
        const int slot = m ? f % o.in_flight : gprt_renderer_next_frame_slot(r);
        // the slot's previous frame is complete once the *_async call below has waited for it; it is written out
        // first so that its host buffer can be reused (the reference's Present happens at the same point)
        if (frame_of[slot] >= 0) {
            if (m) rc = gprt_multi_wait_for_frame_slot(m, slot); // every device's rows of that frame have landed
            else if (gprt_ctx* c = gprt_renderer_frame_context(r, slot)) rc = gprt_sync(c);
            if (rc != GPRT_OK) break;
            if ((rc = write_frame(slot)) != GPRT_OK) break;
        }
        const float t = o.t0 + o.dt * (float)f;
        rc = m ? gprt_multi_on_update(m, t, o.camera_orbit, o.light_orbit) : gprt_renderer_on_update(r, t, o.camera_orbit, o.light_orbit);
        if (rc == GPRT_OK) rc = m ? gprt_multi_on_render_async(m, host[slot]) : gprt_renderer_on_render_async(r, host[slot]);
        frame_of[slot] = f;
    }
    if (rc == GPRT_OK) rc = m ? gprt_multi_wait_for_gpu(m) : gprt_renderer_wait_for_gpu(r);
    const double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
    // the frames still in flight, oldest first
    for (int k = 0; k < o.in_flight && rc == GPRT_OK; k++) {
        int oldest = -1;
        for (int s = 0; s < o.in_flight; s++)
            if (frame_of[s] >= 0 && (oldest < 0 || frame_of[s] < frame_of[oldest])) oldest = s;
        if (oldest < 0) break;
        rc = write_frame(oldest);
        frame_of[oldest] = -1;
    }
    if (rc != GPRT_OK) std::fprintf(stderr, "gprt_render: frame loop failed: %s (%d)\n", gprt_error_string(rc), rc);
    else
        std::printf("scene %d  %ux%u  %zu device context(s)  %d frames  %.3f ms/frame  %.1f fps  ~Million Primary Rays/s: %.1f\n", o.scene, w, h, o.devices.empty() ? (size_t)1 : o.devices.size(), o.frames,
                    seconds * 1e3 / (o.frames - warm), (o.frames - warm) / seconds, (double)w * h * (o.frames - warm) / seconds / 1e6);
    for (void* p : host) gprt_free_host(p);
    if (m) gprt_multi_destroy(m);
    else gprt_renderer_destroy(r);
    return rc == GPRT_OK ? 0 : 1;
}
