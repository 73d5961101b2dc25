// renderer.cpp -- headless Renderer over the gprt C ABI.  See renderer.h.
#include "renderer.h"

#include <cstdio>
#include <cstring>
#include <new>

#include "../../include/gprt_host.h"

namespace gprt_host {

Renderer::Renderer(int scene_id, uint32_t width, uint32_t height, uint32_t grid_n, int device, const gprt_config* config_override)
    : m_device(device) {
    m_valid = build_scene(m_scene, scene_id, width, height, grid_n);
    if (m_valid && config_override) {
        // keep the scene's structural knobs unless the caller overrides them explicitly
        gprt_config c = *config_override;
        if (c.max_recursion_depth == 0) c.max_recursion_depth = m_scene.config.max_recursion_depth;
        if (c.metaballs_count == 0) c.metaballs_count = m_scene.config.metaballs_count;
        if (c.metaball_max_steps == 0) c.metaball_max_steps = m_scene.config.metaball_max_steps;
        if (c.sdf_max_steps == 0) c.sdf_max_steps = m_scene.config.sdf_max_steps;
        if (c.fractal_iterations == 0) c.fractal_iterations = m_scene.config.fractal_iterations;
        c.abi_version = GPRT_ABI_VERSION;
        m_scene.config = c;
    }
    m_device_transforms = m_valid && m_scene.primitives.size() >= 4096;
}

SceneState& Renderer::scene() {
    if (m_host_transforms_stale) {
        update_aabb_primitive_attributes(m_scene, m_scene.animate_geometry_time);
        m_host_transforms_stale = false;
    }
    return m_scene;
}

// What a dispatch passes as per-frame transforms: the host array, or (device mode) nothing after the GPU computed them.
int Renderer::prepare_transforms(gprt_ctx* c, const gprt_prim_per_frame** per_frame, uint32_t* n) {
    if (!m_device_transforms) {
        *per_frame = m_scene.per_frame.data();
        *n = (uint32_t)m_scene.per_frame.size();
        return GPRT_OK;
    }
    *per_frame = nullptr;
    *n = 0;
    return gprt_update_transforms(c, m_scene.animate_geometry_time);
}

Renderer::~Renderer() { on_destroy(); }

void Renderer::on_destroy() {
    for (int i = 1; i < kMaxFrameCount; i++)
        if (m_frame_ctx[i]) { gprt_sync(m_frame_ctx[i]); gprt_destroy(m_frame_ctx[i]); m_frame_ctx[i] = nullptr; }
    m_frame_ctx[0] = nullptr;
    if (m_ctx) {
        gprt_sync(m_ctx);
        gprt_destroy(m_ctx);
        m_ctx = nullptr;
    }
    m_initialized = false;
}

int Renderer::set_frame_count(int n) {
    if (n < 1 || n > kMaxFrameCount) return GPRT_E_INVALID_ARG;
    const int rc = wait_for_gpu();
    if (rc != GPRT_OK) return rc;
    for (int i = n; i < kMaxFrameCount; i++)
        if (i > 0 && m_frame_ctx[i]) { gprt_destroy(m_frame_ctx[i]); m_frame_ctx[i] = nullptr; }
    m_frame_count = n;
    m_frame_index = 0;
    return GPRT_OK;
}

// Frame slots 1.. are created on first use: same geometry, shader table and (deterministic) LBVH.
int Renderer::ensure_frame_contexts() {
    if (!m_initialized) return m_device < 0 ? GPRT_E_NO_DEVICE : GPRT_E_NOT_BUILT;
    m_frame_ctx[0] = m_ctx;
    for (int i = 1; i < m_frame_count; i++) {
        if (m_frame_ctx[i]) continue;
        gprt_ctx* saved = m_ctx;
        int rc = gprt_create(&m_ctx, m_device, &m_scene.config);
        if (rc == GPRT_OK) rc = create_device_dependent_resources();
        gprt_ctx* made = m_ctx;
        m_ctx = saved;
        if (rc != GPRT_OK) { if (made) gprt_destroy(made); return rc; }
        m_frame_ctx[i] = made;
    }
    return GPRT_OK;
}

int Renderer::on_render_async(void* h_rgba8_pinned) {
    int rc = ensure_frame_contexts();
    if (rc != GPRT_OK) return rc;
    gprt_ctx* c = m_frame_ctx[m_frame_index % (uint32_t)m_frame_count];
    m_frame_index++;
    rc = gprt_sync(c); // this slot's previous frame (its staging, queues and host buffer) must be done
    if (rc != GPRT_OK) return rc;
    const gprt_rect tile = {0, 0, m_scene.width, m_scene.height};
    m_last_tile = tile;
    const gprt_prim_per_frame* pf; uint32_t npf;
    rc = prepare_transforms(c, &pf, &npf);
    if (rc != GPRT_OK) return rc;
    return gprt_dispatch_host_async(c, &m_scene.scene_cb, pf, npf, m_scene.width, m_scene.height, tile, h_rgba8_pinned, (size_t)tile.w * 4, nullptr);
}

int Renderer::do_raytracing_async(void* d_rgba8) {
    int rc = ensure_frame_contexts();
    if (rc != GPRT_OK) return rc;
    gprt_ctx* c = m_frame_ctx[m_frame_index % (uint32_t)m_frame_count];
    m_frame_index++;
    rc = gprt_sync(c);
    if (rc != GPRT_OK) return rc;
    const gprt_rect tile = {0, 0, m_scene.width, m_scene.height};
    m_last_tile = tile;
    const gprt_prim_per_frame* pf; uint32_t npf;
    rc = prepare_transforms(c, &pf, &npf);
    if (rc != GPRT_OK) return rc;
    return gprt_dispatch(c, &m_scene.scene_cb, pf, npf, m_scene.width, m_scene.height, tile, d_rgba8, (size_t)tile.w * 4, nullptr, nullptr);
}

int Renderer::do_raytracing_stripes_async(uint32_t stripe_rows, uint32_t rank, uint32_t world, void* d_rgba8) {
    int rc = ensure_frame_contexts();
    if (rc != GPRT_OK) return rc;
    gprt_ctx* c = m_frame_ctx[m_frame_index % (uint32_t)m_frame_count];
    m_frame_index++;
    rc = gprt_sync(c);
    if (rc != GPRT_OK) return rc;
    m_last_tile = {0, 0, m_scene.width, gprt_stripe_local_rows(m_scene.height, stripe_rows, rank, world)};
    const gprt_prim_per_frame* pf; uint32_t npf;
    rc = prepare_transforms(c, &pf, &npf);
    if (rc != GPRT_OK) return rc;
    return gprt_dispatch_stripes(c, &m_scene.scene_cb, pf, npf, m_scene.width, m_scene.height, stripe_rows, rank, world, d_rgba8,
                                 (size_t)m_scene.width * 4, nullptr);
}

int Renderer::on_render_stripes_async(uint32_t stripe_rows, uint32_t rank, uint32_t world, void* h_frame_rgba8) {
    int rc = ensure_frame_contexts();
    if (rc != GPRT_OK) return rc;
    gprt_ctx* c = m_frame_ctx[m_frame_index % (uint32_t)m_frame_count];
    m_frame_index++;
    rc = gprt_sync(c);
    if (rc != GPRT_OK) return rc;
    m_last_tile = {0, 0, m_scene.width, gprt_stripe_local_rows(m_scene.height, stripe_rows, rank, world)};
    const gprt_prim_per_frame* pf; uint32_t npf;
    rc = prepare_transforms(c, &pf, &npf);
    if (rc != GPRT_OK) return rc;
    return gprt_dispatch_stripes_host_async(c, &m_scene.scene_cb, pf, npf, m_scene.width, m_scene.height, stripe_rows, rank, world, h_frame_rgba8,
                                            (size_t)m_scene.width * 4);
}

int Renderer::wait_for_gpu() {
    int rc = GPRT_OK;
    if (m_ctx) rc = gprt_sync(m_ctx);
    for (int i = 1; i < kMaxFrameCount; i++)
        if (m_frame_ctx[i]) { const int r2 = gprt_sync(m_frame_ctx[i]); if (rc == GPRT_OK) rc = r2; }
    return rc;
}

gprt_ctx* Renderer::frame_context(int slot) {
    if (slot < 0 || slot >= m_frame_count) return nullptr;
    return slot == 0 ? m_ctx : m_frame_ctx[slot];
}

int Renderer::on_init() {
    if (!m_valid) return GPRT_E_INVALID_ARG;
    if (m_device < 0) return GPRT_E_NO_DEVICE; // scene-only renderer: there is no CPU render path
    if (m_initialized) return GPRT_OK;
    int rc = gprt_create(&m_ctx, m_device, &m_scene.config);
    if (rc != GPRT_OK) return rc;
    rc = create_device_dependent_resources();
    if (rc != GPRT_OK) { on_destroy(); return rc; }
    m_initialized = true;
    return GPRT_OK;
}

int Renderer::create_device_dependent_resources() {
    int rc = build_geometry();
    if (rc == GPRT_OK) rc = build_shader_tables();
    if (rc == GPRT_OK) rc = build_acceleration_structures();
    if (rc == GPRT_OK) rc = gprt_set_animation(m_ctx, m_scene.anim.data(), (uint32_t)m_scene.anim.size());
    return rc;
}

int Renderer::build_geometry() {
    return gprt_set_geometry(m_ctx, m_scene.aabbs.data(), (uint32_t)m_scene.aabbs.size(), m_scene.vertices.data(),
                             (uint32_t)m_scene.vertices.size(), m_scene.indices.data(), (uint32_t)m_scene.indices.size(),
                             m_scene.instances.data(), (uint32_t)m_scene.instances.size());
}

int Renderer::build_shader_tables() {
    return gprt_set_hit_group_table(m_ctx, m_scene.hit_groups.data(), (uint32_t)m_scene.hit_groups.size(),
                                    (uint32_t)sizeof(gprt_hit_group_record));
}

int Renderer::build_acceleration_structures() { return gprt_build_accel(m_ctx); }

int Renderer::on_update(float animate_geometry_time, float camera_orbit_deg, float light_orbit_deg) {
    if (!m_valid) return GPRT_E_INVALID_ARG;
    if (camera_orbit_deg != 0.0f) rotate_camera_y(m_scene, camera_orbit_deg);
    if (light_orbit_deg != 0.0f) rotate_light_y(m_scene, light_orbit_deg);
    if (m_device_transforms) { set_animation_time(m_scene, animate_geometry_time); m_host_transforms_stale = true; }
    else update_aabb_primitive_attributes(m_scene, animate_geometry_time);
    return GPRT_OK;
}

int Renderer::on_size_changed(uint32_t width, uint32_t height) {
    if (!m_valid || width == 0 || height == 0) return GPRT_E_INVALID_ARG;
    m_scene.width = width;
    m_scene.height = height;
    update_camera_matrices(m_scene);
    return GPRT_OK;
}

static gprt_rect full_if_empty(gprt_rect tile, const SceneState& s) {
    if (tile.w == 0 || tile.h == 0) { tile.x = 0; tile.y = 0; tile.w = s.width; tile.h = s.height; }
    return tile;
}

int Renderer::do_raytracing(gprt_rect tile, void* d_rgba8, uint32_t* d_hit_ids, float* d_rgba_f32) {
    if (!m_initialized) return m_device < 0 ? GPRT_E_NO_DEVICE : GPRT_E_NOT_BUILT;
    tile = full_if_empty(tile, m_scene);
    m_last_tile = tile;
    const gprt_prim_per_frame* pf; uint32_t npf;
    const int rc = prepare_transforms(m_ctx, &pf, &npf);
    if (rc != GPRT_OK) return rc;
    return gprt_dispatch(m_ctx, &m_scene.scene_cb, pf, npf, m_scene.width, m_scene.height, tile, d_rgba8, (size_t)tile.w * 4, d_hit_ids, d_rgba_f32);
}

int Renderer::do_raytracing_stripes(uint32_t stripe_rows, uint32_t rank, uint32_t world, void* d_rgba8, uint32_t* d_hit_ids) {
    if (!m_initialized) return m_device < 0 ? GPRT_E_NO_DEVICE : GPRT_E_NOT_BUILT;
    m_last_tile = {0, 0, m_scene.width, gprt_stripe_local_rows(m_scene.height, stripe_rows, rank, world)};
    const gprt_prim_per_frame* pf; uint32_t npf;
    const int rc = prepare_transforms(m_ctx, &pf, &npf);
    if (rc != GPRT_OK) return rc;
    return gprt_dispatch_stripes(m_ctx, &m_scene.scene_cb, pf, npf, m_scene.width, m_scene.height, stripe_rows, rank, world, d_rgba8,
                                 (size_t)m_scene.width * 4, d_hit_ids);
}

int Renderer::on_render(gprt_rect tile, void* h_rgba8, uint32_t* h_hit_ids) {
    if (!m_initialized) return m_device < 0 ? GPRT_E_NO_DEVICE : GPRT_E_NOT_BUILT;
    tile = full_if_empty(tile, m_scene);
    m_last_tile = tile;
    const gprt_prim_per_frame* pf; uint32_t npf;
    const int rc = prepare_transforms(m_ctx, &pf, &npf);
    if (rc != GPRT_OK) return rc;
    return gprt_dispatch_host(m_ctx, &m_scene.scene_cb, pf, npf, m_scene.width, m_scene.height, tile, h_rgba8, (size_t)tile.w * 4, h_hit_ids);
}

float Renderer::primary_mrays_per_s() {
    if (!m_initialized) return 0.0f;
    gprt_timing t;
    if (gprt_last_timing(m_ctx, &t) != GPRT_OK || t.dispatch_ms <= 0.0f) return 0.0f;
    return (float)((double)m_last_tile.w * m_last_tile.h / (t.dispatch_ms * 1e-3) / 1e6);
}

} // namespace gprt_host

// ---------------------------------------------------------------------------------------------
// C wrappers
// ---------------------------------------------------------------------------------------------
struct gprt_renderer {
    gprt_host::Renderer impl;
    gprt_renderer(int scene_id, uint32_t w, uint32_t h, uint32_t grid_n, int device, const gprt_config* cfg)
        : impl(scene_id, w, h, grid_n, device, cfg) {}
};

extern "C" {

int gprt_renderer_create(gprt_renderer** out, int scene_id, uint32_t width, uint32_t height, uint32_t grid_n, int device,
                         const gprt_config* config_override) {
    if (!out) return GPRT_E_INVALID_ARG;
    *out = nullptr;
    gprt_renderer* r = new (std::nothrow) gprt_renderer(scene_id, width, height, grid_n, device, config_override);
    if (!r) return GPRT_E_OUT_OF_MEMORY;
    if (!r->impl.valid()) { delete r; return GPRT_E_INVALID_ARG; }
    *out = r;
    return GPRT_OK;
}
void gprt_renderer_destroy(gprt_renderer* r) { delete r; }
int gprt_renderer_on_init(gprt_renderer* r) { return r ? r->impl.on_init() : GPRT_E_INVALID_ARG; }
int gprt_renderer_on_update(gprt_renderer* r, float t, float cam_deg, float light_deg) {
    return r ? r->impl.on_update(t, cam_deg, light_deg) : GPRT_E_INVALID_ARG;
}
int gprt_renderer_on_render(gprt_renderer* r, gprt_rect tile, void* h_rgba8, uint32_t* h_hit_ids) {
    return r ? r->impl.on_render(tile, h_rgba8, h_hit_ids) : GPRT_E_INVALID_ARG;
}
int gprt_renderer_do_raytracing(gprt_renderer* r, gprt_rect tile, void* d_rgba8, uint32_t* d_hit_ids, float* d_rgba_f32) {
    return r ? r->impl.do_raytracing(tile, d_rgba8, d_hit_ids, d_rgba_f32) : GPRT_E_INVALID_ARG;
}
int gprt_renderer_do_raytracing_stripes(gprt_renderer* r, uint32_t stripe_rows, uint32_t rank, uint32_t world, void* d_rgba8,
                                        uint32_t* d_hit_ids) {
    return r ? r->impl.do_raytracing_stripes(stripe_rows, rank, world, d_rgba8, d_hit_ids) : GPRT_E_INVALID_ARG;
}
int gprt_renderer_on_render_async(gprt_renderer* r, void* h_rgba8_pinned) { return r ? r->impl.on_render_async(h_rgba8_pinned) : GPRT_E_INVALID_ARG; }
int gprt_renderer_do_raytracing_async(gprt_renderer* r, void* d_rgba8) { return r ? r->impl.do_raytracing_async(d_rgba8) : GPRT_E_INVALID_ARG; }
int gprt_renderer_do_raytracing_stripes_async(gprt_renderer* r, uint32_t stripe_rows, uint32_t rank, uint32_t world, void* d_rgba8) {
    return r ? r->impl.do_raytracing_stripes_async(stripe_rows, rank, world, d_rgba8) : GPRT_E_INVALID_ARG;
}
int gprt_renderer_on_render_stripes_async(gprt_renderer* r, uint32_t stripe_rows, uint32_t rank, uint32_t world, void* h_frame_rgba8) {
    return r ? r->impl.on_render_stripes_async(stripe_rows, rank, world, h_frame_rgba8) : GPRT_E_INVALID_ARG;
}
int gprt_renderer_get_device_transforms(gprt_renderer* r) { return r && r->impl.device_transforms() ? 1 : 0; }
int gprt_renderer_set_device_transforms(gprt_renderer* r, int on) { if (!r) return GPRT_E_INVALID_ARG; r->impl.set_device_transforms(on != 0); return GPRT_OK; }
int gprt_renderer_create_frame_contexts(gprt_renderer* r) { return r ? r->impl.create_frame_contexts() : GPRT_E_INVALID_ARG; }
int gprt_renderer_set_frame_count(gprt_renderer* r, int n) { return r ? r->impl.set_frame_count(n) : GPRT_E_INVALID_ARG; }
int gprt_renderer_get_frame_count(gprt_renderer* r) { return r ? r->impl.frame_count() : -1; }
int gprt_renderer_next_frame_slot(gprt_renderer* r) { return r ? r->impl.next_frame_slot() : -1; }
int gprt_renderer_wait_for_gpu(gprt_renderer* r) { return r ? r->impl.wait_for_gpu() : GPRT_E_INVALID_ARG; }
gprt_ctx* gprt_renderer_frame_context(gprt_renderer* r, int slot) { return r ? r->impl.frame_context(slot) : nullptr; }
int gprt_renderer_on_size_changed(gprt_renderer* r, uint32_t w, uint32_t h) {
    return r ? r->impl.on_size_changed(w, h) : GPRT_E_INVALID_ARG;
}
gprt_ctx* gprt_renderer_context(gprt_renderer* r) { return r ? r->impl.context() : nullptr; }
float gprt_renderer_primary_mrays_per_s(gprt_renderer* r) { return r ? r->impl.primary_mrays_per_s() : 0.0f; }

int gprt_renderer_get_size(gprt_renderer* r, uint32_t* width, uint32_t* height) {
    if (!r || !width || !height) return GPRT_E_INVALID_ARG;
    *width = r->impl.size_w(); *height = r->impl.size_h();
    return GPRT_OK;
}

int gprt_renderer_scene(gprt_renderer* r, gprt_scene_view* out) {
    if (!r || !out) return GPRT_E_INVALID_ARG;
    gprt_host::SceneState& s = r->impl.scene();
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

int gprt_write_ppm(const char* path, const void* rgba8, uint32_t width, uint32_t height) {
    if (!path || !rgba8) return GPRT_E_INVALID_ARG;
    FILE* f = std::fopen(path, "wb");
    if (!f) return GPRT_E_INVALID_ARG;
    std::fprintf(f, "P6\n%u %u\n255\n", width, height);
    const unsigned char* p = static_cast<const unsigned char*>(rgba8);
    for (size_t i = 0; i < (size_t)width * height; i++) std::fwrite(p + 4 * i, 1, 3, f);
    std::fclose(f);
    return GPRT_OK;
}

} // extern "C"
