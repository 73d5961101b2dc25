/*
 * gprt_host.h -- C wrappers over the headless C++ Renderer (gpuraytracer_b200/host/renderer.h),
 * the host side that replaces the reference's Renderer / DeviceResources / Window trio
 * (src/Renderer.h:24-28 on_init / on_update / on_render / on_size_changed / on_destroy).
 * Used by the Python bindings (tests, bench.py) and the gprt_render CLI.
 */
#ifndef GPRT_HOST_H
#define GPRT_HOST_H

#include "gprt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gprt_renderer gprt_renderer;

/* Read-only view of the scene state the renderer would hand to gprt_set_geometry /
 * gprt_set_hit_group_table / gprt_dispatch.  Pointers stay valid until the next call on the
 * renderer. */
typedef struct gprt_scene_view {
    uint32_t width, height;
    gprt_config config;
    const gprt_scene_cb* scene_cb;
    const gprt_prim_per_frame* per_frame;
    uint32_t n_per_frame;
    const gprt_aabb* aabbs;
    uint32_t n_aabbs;
    const gprt_hit_group_record* hit_groups;
    uint32_t n_hit_groups;
    const gprt_vertex* vertices;
    uint32_t n_vertices;
    const uint16_t* indices;
    uint32_t n_indices;
    const gprt_instance_desc* instances;
    uint32_t n_instances;
} gprt_scene_view;

/* scene_id: 0 = reference scene, 1..5 = BASELINE configs, 6 = config 3 with the long march
 * budget.  width/height 0 = the config's own size.  grid_n: config-5 grid edge (0 = 1000).
 * device < 0 builds the scene only (no CUDA context; on_render then fails with
 * GPRT_E_NO_DEVICE) -- that is how CPU-only tests obtain scene data for the oracle. */
int gprt_renderer_create(gprt_renderer** out, int scene_id, uint32_t width, uint32_t height,
                         uint32_t grid_n, int device, const gprt_config* config_override);
void gprt_renderer_destroy(gprt_renderer* r);

/* Renderer::on_init: device objects, geometry upload, acceleration structures. */
int gprt_renderer_on_init(gprt_renderer* r);
/* Renderer::on_update with the StepTimer replaced by an explicit animation time; the optional
 * camera/light orbits (src/Renderer.cpp:92-110) take degrees for this update. */
int gprt_renderer_on_update(gprt_renderer* r, float animate_geometry_time,
                            float camera_orbit_deg, float light_orbit_deg);
/* Renderer::on_render: do_raytracing over `tile` (w == 0: whole image) and, instead of the copy
 * to the swap chain, an optional copy to host memory (tile-packed RGBA8 / hit ids). */
int gprt_renderer_on_render(gprt_renderer* r, gprt_rect tile, void* h_rgba8, uint32_t* h_hit_ids);
/* do_raytracing only, output left in device memory (tile-packed); asynchronous. */
int gprt_renderer_do_raytracing(gprt_renderer* r, gprt_rect tile, void* d_rgba8, uint32_t* d_hit_ids,
                                float* d_rgba_f32);
/* Multi-GPU: do_raytracing over this rank's row stripes only (gprt_dispatch_stripes); the output is
 * gprt_stripe_local_rows() packed rows of width * 4 bytes in device memory; asynchronous. */
int gprt_renderer_do_raytracing_stripes(gprt_renderer* r, uint32_t stripe_rows, uint32_t rank, uint32_t world,
                                        void* d_rgba8, uint32_t* d_hit_ids);
/* Frames in flight (frame_count = 3, src/Renderer.h:92): frame k is rendered by context k % 3 on its
 * own stream; each call first waits for that slot's previous frame.  h_rgba8_pinned / d_rgba8 must
 * stay valid until gprt_renderer_wait_for_gpu (or until the slot comes round again). */
int gprt_renderer_on_render_async(gprt_renderer* r, void* h_rgba8_pinned);
int gprt_renderer_do_raytracing_async(gprt_renderer* r, void* d_rgba8);
int gprt_renderer_do_raytracing_stripes_async(gprt_renderer* r, uint32_t stripe_rows, uint32_t rank, uint32_t world, void* d_rgba8);
/* Multi-GPU with host output (gprt_dispatch_stripes_host_async): this rank's rows are copied to their global rows of
 * h_frame_rgba8, a tightly pitched width * height * 4 host frame shared by all ranks (pinned / registered). */
int gprt_renderer_on_render_stripes_async(gprt_renderer* r, uint32_t stripe_rows, uint32_t rank, uint32_t world, void* h_frame_rgba8);
/* Compute the per-frame transforms on the GPU (gprt_update_transforms) instead of on the host + 128 B/primitive upload.
 * Default: on for scenes of >= 4096 primitives. */
int gprt_renderer_set_device_transforms(gprt_renderer* r, int on);
int gprt_renderer_get_device_transforms(gprt_renderer* r);
/* Frames in flight: 3 like the reference (src/Renderer.h:92) unless set (1..8); waits for the GPU, restarts at slot 0. */
int gprt_renderer_set_frame_count(gprt_renderer* r, int n);
int gprt_renderer_get_frame_count(gprt_renderer* r);
int gprt_renderer_create_frame_contexts(gprt_renderer* r); /* create the frame contexts (and their streams) up front */
int gprt_renderer_next_frame_slot(gprt_renderer* r);       /* slot the next *_async call will use */
int gprt_renderer_wait_for_gpu(gprt_renderer* r);
gprt_ctx* gprt_renderer_frame_context(gprt_renderer* r, int slot); /* NULL until the slot was used */
int gprt_renderer_on_size_changed(gprt_renderer* r, uint32_t width, uint32_t height);

int gprt_renderer_scene(gprt_renderer* r, gprt_scene_view* out); /* refreshes the host transforms if they are computed on the device */
int gprt_renderer_get_size(gprt_renderer* r, uint32_t* width, uint32_t* height);
gprt_ctx* gprt_renderer_context(gprt_renderer* r);

/* Frame statistics in the reference's convention (calculate_frame_stats,
 * src/Renderer.cpp:375-399): million PRIMARY rays per second from the last GPU dispatch time. */
float gprt_renderer_primary_mrays_per_s(gprt_renderer* r);

/* One frame cut over several CUDA devices from ONE process (host/multi_renderer.h): entry k of `devices` renders the row
 * stripes k, k + N, ... of every frame with its own context (replicated scene) and copies its rows to their global rows of
 * the caller's frame -- a page-locked host frame (every GPU over its own PCIe link) or a device frame on devices[0] (peer
 * copies over NVLink).  A device may appear more than once: {0, 0, 0} renders one frame as three concurrent stripe
 * dispatches on one GPU, overlapping their serial tails. */
typedef struct gprt_multi gprt_multi;
int gprt_multi_create(gprt_multi** out, int scene_id, uint32_t width, uint32_t height, uint32_t grid_n,
                      const int* devices, int n_devices, const gprt_config* config_override);
void gprt_multi_destroy(gprt_multi* m);
int gprt_multi_on_init(gprt_multi* m);
int gprt_multi_on_update(gprt_multi* m, float animate_geometry_time, float camera_orbit_deg, float light_orbit_deg);
int gprt_multi_on_render_async(gprt_multi* m, void* h_frame_rgba8);     /* tightly pitched width * height * 4, page-locked */
int gprt_multi_do_raytracing_async(gprt_multi* m, void* d_frame_rgba8); /* device memory of devices[0]                      */
int gprt_multi_wait_for_gpu(gprt_multi* m);
int gprt_multi_wait_for_frame_slot(gprt_multi* m, int slot);            /* frame k runs on slot k % frames-in-flight of every device */
int gprt_multi_set_frame_count(gprt_multi* m, int n);                   /* frames in flight per device (1..8)               */
int gprt_multi_get_size(gprt_multi* m, uint32_t* width, uint32_t* height);
uint32_t gprt_multi_stripe_rows(gprt_multi* m);
gprt_ctx* gprt_multi_context(gprt_multi* m, int i);                     /* frame slot 0 of entry i                          */

/* The CUDA-free scene builder on its own (also exported by libgprt_scene.so, which links no CUDA): the arrays
 * Renderer::initialize_scene / build_geometry / build_shader_tables assemble on the host
 * (src/Renderer.cpp:197-356, 478-565, 732-786, 839-985) for scene_id at animation time t. */
typedef struct gprt_scene gprt_scene;
int gprt_scene_create(gprt_scene** out, int scene_id, uint32_t width, uint32_t height, uint32_t grid_n);
void gprt_scene_destroy(gprt_scene* s);
int gprt_scene_update(gprt_scene* s, float animate_geometry_time, float camera_orbit_deg, float light_orbit_deg);
int gprt_scene_get(gprt_scene* s, gprt_scene_view* out); /* pointers stay valid until the next call on s */

/* Write a tile-packed RGBA8 buffer as a binary PPM (the headless stand-in for Present). */
int gprt_write_ppm(const char* path, const void* rgba8, uint32_t width, uint32_t height);

#ifdef __cplusplus
}
#endif
#endif /* GPRT_HOST_H */
