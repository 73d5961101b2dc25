// scene.h -- host-side scene state of the headless Renderer (C++).
//
// Mirrors the data members and the scene-building members of the reference's Renderer
// (src/Renderer.h:101-127; src/Renderer.cpp:197-356 initialize_scene / update_camera_matrices /
// update_aabb_primitive_attributes, :478-565 build_geometry, :732-786 instance descs,
// :839-985 build_shader_tables), minus everything D3D12.  The DirectXMath calls the reference
// makes are restated in mathx.h (row-vector convention, row-major storage).
#pragma once
#include <cstdint>
#include <vector>

#include "../../include/gprt.h"

namespace gprt_host {

// One procedural primitive of the scene: what the reference spreads over m_aabbs[i],
// m_aabb_material_cb[i], the hit-group record i and set_transform_for_aabb(i, ...).
struct PrimitiveSpec {
    uint32_t hit_group_kind;  // GPRT_HG_AABB_*
    uint32_t primitive_type;  // enum inside the class
    gprt_material material;
    float scale[3];           // XMMatrixScaling
    bool rotates;             // uses XMMatrixRotationY(-2 t)
    gprt_aabb aabb;           // BLAS object space
};

struct SceneState {
    uint32_t width = 1280, height = 720; // src/main.cpp:14
    gprt_config config;

    std::vector<PrimitiveSpec> primitives;
    gprt_material plane_material;
    float plane_extent[3];     // XMMatrixScaling of the quad instance
    float plane_translate[3];
    float aabb_instance_translate[3];

    // camera (src/Renderer.h:108-110)
    double eye[4], at[4], up[4];
    float fov_angle_y_deg = 45.0f, z_near = 0.01f, z_far = 125.0f;

    // what do_raytracing uploads / binds
    gprt_scene_cb scene_cb;
    std::vector<gprt_prim_per_frame> per_frame;
    std::vector<gprt_aabb> aabbs;
    std::vector<gprt_prim_anim> anim;   // static inputs of update_aabb_primitive_attributes (for the on-device update)
    std::vector<gprt_hit_group_record> hit_groups;
    std::vector<gprt_vertex> vertices;
    std::vector<uint16_t> indices;
    std::vector<gprt_instance_desc> instances; // the TLAS: [0] the ground quad, [1] the AABB BLAS (src/Renderer.cpp:732-786); kInstanced adds more
    bool extra_instances = false;              // kInstanced: three AABB instances + a second triangle instance

    float animate_geometry_time = 0.0f;
};

// Scene ids: 0 = the reference scene at the caller's size (src/Renderer.cpp);
// 1..5 = the five BASELINE.json configs made concrete.
// 7 = the reference primitives under a TLAS with several instances of each BLAS (two-level instancing).
enum SceneId { kSceneReference = 0, kConfig1 = 1, kConfig2 = 2, kConfig3 = 3, kConfig4 = 4, kConfig5 = 5, kConfig3Long = 6, kInstanced = 7 };

// width/height == 0 selects the config's own resolution.  grid_n overrides the grid edge of
// config 5 (0 = 1000) so tests can use a small instance grid.
bool build_scene(SceneState& s, int scene_id, uint32_t width, uint32_t height, uint32_t grid_n);

void initialize_camera(SceneState& s, const double eye[3], const double at[3]); // initialize_scene, camera part
void update_camera_matrices(SceneState& s);                                     // src/Renderer.cpp:289-300
void update_aabb_primitive_attributes(SceneState& s, float animation_time);     // src/Renderer.cpp:302-356
void set_animation_time(SceneState& s, float animation_time);                   // the same without the per-primitive matrices
void rotate_camera_y(SceneState& s, float degrees);                             // on_update camera orbit, :92-101
void rotate_light_y(SceneState& s, float degrees);                              // on_update light orbit, :104-110
void build_geometry_and_tables(SceneState& s);                                  // build_geometry + instance descs + shader tables

} // namespace gprt_host
