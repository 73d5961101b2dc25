// multi_renderer.h -- one frame cut over several CUDA devices of the box from ONE process (the reference
// renders on one adapter, src/DeviceResources.cpp): every entry of `devices` gets its own Renderer (replicated scene, its
// own deterministic LBVH, its own frames in flight) and renders the interleaved row stripes r, r + N, ... of every frame
// (gprt_dispatch_stripes: DispatchRaysIndex / Dimensions stay global, so the union is bit-identical to one dispatch).
// Each device copies its rows straight to their global rows of the caller's frame:
//   on_render_async      a page-locked HOST frame shared by all devices -- every GPU delivers over its own PCIe link;
//   do_raytracing_async  a DEVICE frame on devices[0] -- peer copies over NVLink (the composite of the north star).
// A device may be listed more than once: the entries are independent contexts, so {0, 0, 0} renders ONE frame as three
// concurrent stripe dispatches on one GPU -- the serial tail of each (a few long rays per wave) overlaps the bulk of the
// others inside a single frame, which is what frames in flight do across frames.
#pragma once
#include <memory>
#include <vector>

#include "renderer.h"

namespace gprt_host {

class MultiRenderer {
public:
    MultiRenderer(int scene_id, uint32_t width, uint32_t height, uint32_t grid_n, const int* devices, int n_devices, const gprt_config* config_override);
    ~MultiRenderer();
    bool valid() const { return m_valid; }
    int on_init();
    int on_update(float animate_geometry_time, float camera_orbit_deg, float light_orbit_deg);
    int on_render_async(void* h_frame_rgba8);        // host frame, width * 4 bytes per row, page-locked
    int do_raytracing_async(void* d_frame_rgba8);    // device frame on devices[0]
    int wait_for_gpu();
    int wait_for_frame_slot(int slot);               // the frame last issued on frame slot `slot` has landed from every device
    int set_frame_count(int n);
    void on_destroy();
    uint32_t stripe_rows() const { return m_stripe_rows; }
    uint32_t size_w() const { return m_r.empty() ? 0 : m_r[0]->size_w(); }
    uint32_t size_h() const { return m_r.empty() ? 0 : m_r[0]->size_h(); }
    int n_devices() const { return (int)m_r.size(); }
    Renderer* renderer(int i) { return (i >= 0 && i < (int)m_r.size()) ? m_r[i].get() : nullptr; }

private:
    std::vector<std::unique_ptr<Renderer>> m_r;
    std::vector<int> m_devices;
    uint32_t m_stripe_rows = 1;
    bool m_valid = false;
};

} // namespace gprt_host
