// multi_renderer.cpp -- see multi_renderer.h.
#include "multi_renderer.h"

#include <new>

#include "../../include/gprt_host.h"

namespace gprt_host {

namespace {
// Largest stripe height <= 8 rows that gives every device the same number of whole stripes; 1 if there is none.  A few
// rows per stripe keep neighbouring rays together; interleaving spreads the expensive image regions over all devices.
uint32_t stripe_rows_for(uint32_t height, uint32_t world) {
    if (world <= 1) return height ? height : 1;
    for (uint32_t s = 8; s >= 1; s--)
        if (height % (s * world) == 0) return s;
    return 1;
}
} // namespace

MultiRenderer::MultiRenderer(int scene_id, uint32_t width, uint32_t height, uint32_t grid_n, const int* devices, int n_devices,
                             const gprt_config* config_override) {
    if (!devices || n_devices < 1 || n_devices > 64) return;
    for (int i = 0; i < n_devices; i++) {
        m_devices.push_back(devices[i]);
        m_r.emplace_back(new (std::nothrow) Renderer(scene_id, width, height, grid_n, devices[i], config_override));
        if (!m_r.back() || !m_r.back()->valid()) return;
    }
    m_stripe_rows = stripe_rows_for(m_r[0]->size_h(), (uint32_t)n_devices);
    m_valid = true;
}

MultiRenderer::~MultiRenderer() { on_destroy(); }

void MultiRenderer::on_destroy() {
    for (auto& r : m_r)
        if (r) r->on_destroy();
}

int MultiRenderer::on_init() {
    if (!m_valid) return GPRT_E_INVALID_ARG;
    for (auto& r : m_r) {
        const int rc = r->on_init();
        if (rc != GPRT_OK) return rc;
    }
    // peer access towards devices[0], where the device-resident composite lives (ignored where it is not available:
    // the copies then go through the host)
    for (size_t i = 1; i < m_r.size(); i++)
        if (m_devices[i] != m_devices[0]) gprt_enable_peer_access(m_r[i]->context(), m_devices[0]);
    return GPRT_OK;
}

int MultiRenderer::set_frame_count(int n) {
    for (auto& r : m_r) {
        const int rc = r->set_frame_count(n);
        if (rc != GPRT_OK) return rc;
    }
    return GPRT_OK;
}

int MultiRenderer::on_update(float t, float camera_orbit_deg, float light_orbit_deg) {
    for (auto& r : m_r) {
        const int rc = r->on_update(t, camera_orbit_deg, light_orbit_deg);
        if (rc != GPRT_OK) return rc;
    }
    return GPRT_OK;
}

int MultiRenderer::on_render_async(void* h_frame_rgba8) {
    if (!m_valid || !h_frame_rgba8) return GPRT_E_INVALID_ARG;
    const uint32_t world = (uint32_t)m_r.size();
    for (uint32_t k = 0; k < world; k++) {
        const int rc = m_r[k]->on_render_stripes_async(m_stripe_rows, k, world, h_frame_rgba8);
        if (rc != GPRT_OK) return rc;
    }
    return GPRT_OK;
}

int MultiRenderer::do_raytracing_async(void* d_frame_rgba8) { return on_render_async(d_frame_rgba8); } // the copies are cudaMemcpyDefault: host or peer

int MultiRenderer::wait_for_gpu() {
    int rc = GPRT_OK;
    for (auto& r : m_r) {
        const int r2 = r->wait_for_gpu();
        if (rc == GPRT_OK) rc = r2;
    }
    return rc;
}

int MultiRenderer::wait_for_frame_slot(int slot) {
    int rc = GPRT_OK;
    for (auto& r : m_r) {
        gprt_ctx* c = r->frame_context(slot);
        const int r2 = c ? gprt_sync(c) : GPRT_OK;
        if (rc == GPRT_OK) rc = r2;
    }
    return rc;
}

} // namespace gprt_host

// ---------------------------------------------------------------------------------------------
// C wrappers
// ---------------------------------------------------------------------------------------------
struct gprt_multi {
    gprt_host::MultiRenderer impl;
    gprt_multi(int scene_id, uint32_t w, uint32_t h, uint32_t grid_n, const int* devices, int n, const gprt_config* cfg)
        : impl(scene_id, w, h, grid_n, devices, n, cfg) {}
};

extern "C" {

int gprt_multi_create(gprt_multi** out, int scene_id, uint32_t width, uint32_t height, uint32_t grid_n, const int* devices, int n_devices,
                      const gprt_config* config_override) {
    if (!out) return GPRT_E_INVALID_ARG;
    *out = nullptr;
    gprt_multi* m = new (std::nothrow) gprt_multi(scene_id, width, height, grid_n, devices, n_devices, config_override);
    if (!m) return GPRT_E_OUT_OF_MEMORY;
    if (!m->impl.valid()) { delete m; return GPRT_E_INVALID_ARG; }
    *out = m;
    return GPRT_OK;
}
void gprt_multi_destroy(gprt_multi* m) { delete m; }
int gprt_multi_on_init(gprt_multi* m) { return m ? m->impl.on_init() : GPRT_E_INVALID_ARG; }
int gprt_multi_on_update(gprt_multi* m, float t, float cam_deg, float light_deg) { return m ? m->impl.on_update(t, cam_deg, light_deg) : GPRT_E_INVALID_ARG; }
int gprt_multi_on_render_async(gprt_multi* m, void* h_frame_rgba8) { return m ? m->impl.on_render_async(h_frame_rgba8) : GPRT_E_INVALID_ARG; }
int gprt_multi_do_raytracing_async(gprt_multi* m, void* d_frame_rgba8) { return m ? m->impl.do_raytracing_async(d_frame_rgba8) : GPRT_E_INVALID_ARG; }
int gprt_multi_wait_for_frame_slot(gprt_multi* m, int slot) { return m ? m->impl.wait_for_frame_slot(slot) : GPRT_E_INVALID_ARG; }
int gprt_multi_wait_for_gpu(gprt_multi* m) { return m ? m->impl.wait_for_gpu() : GPRT_E_INVALID_ARG; }
int gprt_multi_set_frame_count(gprt_multi* m, int n) { return m ? m->impl.set_frame_count(n) : GPRT_E_INVALID_ARG; }
int gprt_multi_get_size(gprt_multi* m, uint32_t* width, uint32_t* height) {
    if (!m || !width || !height) return GPRT_E_INVALID_ARG;
    *width = m->impl.size_w(); *height = m->impl.size_h();
    return GPRT_OK;
}
uint32_t gprt_multi_stripe_rows(gprt_multi* m) { return m ? m->impl.stripe_rows() : 0u; }
gprt_ctx* gprt_multi_context(gprt_multi* m, int i) {
    gprt_host::Renderer* r = m ? m->impl.renderer(i) : nullptr;
    return r ? r->context() : nullptr;
}

} // extern "C"
