// scene_capi.cpp -- C wrappers over the CUDA-free scene builder (scene.cpp): the host-side data the
// reference's Renderer assembles before it ever touches the device (src/Renderer.cpp:197-356,
// 478-565, 732-786, 839-985).  Compiled into libgprt.so and, on its own with g++, into
// libgprt_scene.so, which has no CUDA dependency: the bench's reference arm and CPU-only tools
// obtain scene constants from it without mapping the product library.
#include <new>

#include "../../include/gprt_host.h"
#include "scene.h"

struct gprt_scene {
    gprt_host::SceneState s;
};

extern "C" {

int gprt_scene_create(gprt_scene** out, int scene_id, uint32_t width, uint32_t height, uint32_t grid_n) {
    if (!out) return GPRT_E_INVALID_ARG;
    *out = nullptr;
    gprt_scene* sc = new (std::nothrow) gprt_scene();
    if (!sc) return GPRT_E_OUT_OF_MEMORY;
    if (!gprt_host::build_scene(sc->s, scene_id, width, height, grid_n)) { delete sc; return GPRT_E_INVALID_ARG; }
    *out = sc;
    return GPRT_OK;
}

void gprt_scene_destroy(gprt_scene* sc) { delete sc; }

int gprt_scene_update(gprt_scene* sc, float animate_geometry_time, float camera_orbit_deg, float light_orbit_deg) {
    if (!sc) return GPRT_E_INVALID_ARG;
    if (camera_orbit_deg != 0.0f) gprt_host::rotate_camera_y(sc->s, camera_orbit_deg);
    if (light_orbit_deg != 0.0f) gprt_host::rotate_light_y(sc->s, light_orbit_deg);
    gprt_host::update_aabb_primitive_attributes(sc->s, animate_geometry_time);
    return GPRT_OK;
}

int gprt_scene_get(gprt_scene* sc, gprt_scene_view* out) {
    if (!sc || !out) return GPRT_E_INVALID_ARG;
    gprt_host::SceneState& s = sc->s;
    out->width = s.width; out->height = s.height;
    out->config = s.config;
    out->scene_cb = &s.scene_cb;
    out->per_frame = s.per_frame.data(); out->n_per_frame = (uint32_t)s.per_frame.size();
    out->aabbs = s.aabbs.data(); out->n_aabbs = (uint32_t)s.aabbs.size();
    out->hit_groups = s.hit_groups.data(); out->n_hit_groups = (uint32_t)s.hit_groups.size();
    out->vertices = s.vertices.data(); out->n_vertices = (uint32_t)s.vertices.size();
    out->indices = s.indices.data(); out->n_indices = (uint32_t)s.indices.size();
    out->instances = s.instances.data(); out->n_instances = (uint32_t)s.instances.size();
    return GPRT_OK;
}

} // extern "C"
