// renderer.h -- headless C++ Renderer: same member names and call order as the reference's
// Renderer (src/Renderer.h:15-135), with the D3D12 device, swap chain and window replaced by a
// gprt_ctx (CUDA) and a host/device framebuffer.
#pragma once
#include <cstdint>

#include "../../include/gprt.h"
#include "scene.h"

namespace gprt_host {

class Renderer {
public:
    Renderer(int scene_id, uint32_t width, uint32_t height, uint32_t grid_n, int device, const gprt_config* config_override);
    ~Renderer();

    int on_init();                                               // src/Renderer.cpp:56-80
    int on_update(float animate_geometry_time, float camera_orbit_deg, float light_orbit_deg); // :82-120
    int on_render(gprt_rect tile, void* h_rgba8, uint32_t* h_hit_ids); // :122-148
    int on_size_changed(uint32_t width, uint32_t height);        // :150-161
    void on_destroy();                                           // :163-170

    int do_raytracing(gprt_rect tile, void* d_rgba8, uint32_t* d_hit_ids, float* d_rgba_f32); // :401-452
    // multi-GPU: this rank's row stripes only (gprt_dispatch_stripes), packed output
    int do_raytracing_stripes(uint32_t stripe_rows, uint32_t rank, uint32_t world, void* d_rgba8, uint32_t* d_hit_ids);
    // Frames in flight (src/Renderer.h:92 frame_count = 3, DeviceResources::move_to_next_frame): frame k
    // is rendered by context k % 3 on that context's own stream, so the tail of one frame (a few long
    // rays) overlaps the bulk of the next and the copy to the host overlaps rendering.  Each call
    // first waits for the slot's previous frame, like the reference waits on its frame fence.
    int on_render_async(void* h_rgba8_pinned);                   // on_render, 3 frames in flight, whole image
    int do_raytracing_async(void* d_rgba8);                      // same with device output
    int do_raytracing_stripes_async(uint32_t stripe_rows, uint32_t rank, uint32_t world, void* d_rgba8); // multi-GPU, 3 in flight
    // multi-GPU, host output: this rank's rows go straight to their global rows of a full-frame host buffer shared by all ranks
    int on_render_stripes_async(uint32_t stripe_rows, uint32_t rank, uint32_t world, void* h_frame_rgba8);
    int create_frame_contexts() { return ensure_frame_contexts(); }
    int next_frame_slot() const { return (int)(m_frame_index % (uint32_t)m_frame_count); }
    // Frames in flight: 3 like the reference unless changed (1..kMaxFrameCount).  More slots hide the serial tail of a
    // frame better when a GPU renders only a stripe of it; each slot owns a full set of wavefront queues.
    int set_frame_count(int n);
    int frame_count() const { return m_frame_count; }
    int wait_for_gpu();                                          // DeviceResources::wait_for_gpu: all frames done
    gprt_ctx* frame_context(int slot);
    float primary_mrays_per_s();                                 // calculate_frame_stats, :375-399

    // Per-frame transforms on the GPU (gprt_update_transforms) instead of host arithmetic + upload; on by default for
    // scenes of >= 4096 primitives.  The host copy is then refreshed only when scene() is asked for it.
    void set_device_transforms(bool on) { m_device_transforms = on; }
    bool device_transforms() const { return m_device_transforms; }
    SceneState& scene();
    uint32_t size_w() const { return m_scene.width; }
    uint32_t size_h() const { return m_scene.height; }
    gprt_ctx* context() { return m_ctx; }
    bool valid() const { return m_valid; }

private:
    int create_device_dependent_resources();  // :987-1021
    int build_geometry();                      // :478-565  -> gprt_set_geometry
    int build_shader_tables();                 // :839-985  -> gprt_set_hit_group_table
    int build_acceleration_structures();       // :788-835  -> gprt_build_accel

    SceneState m_scene;
    static constexpr int kFrameCount = 3;     // src/Renderer.h:92
    static constexpr int kMaxFrameCount = 8;
    int m_frame_count = kFrameCount;
    int ensure_frame_contexts();
    gprt_ctx* m_ctx = nullptr;           // frame slot 0 (and the synchronous path)
    gprt_ctx* m_frame_ctx[kMaxFrameCount] = {};
    uint32_t m_frame_index = 0;
    int m_device = -1;
    int prepare_transforms(gprt_ctx* c, const gprt_prim_per_frame** per_frame, uint32_t* n);
    bool m_device_transforms = false;
    bool m_host_transforms_stale = false;
    bool m_valid = false;
    bool m_initialized = false;
    gprt_rect m_last_tile = {0, 0, 0, 0};
};

} // namespace gprt_host
