// scene.cpp -- scene construction of the headless Renderer.  See scene.h for the mapping to
// the reference's Renderer members.  All constants below are the reference's
// (SURVEY.md Appendix A cites each one); the five BASELINE configs are synthetic
// arrangements of the same ten primitive templates.
#include "scene.h"

#include <cstring>

#include "mathx.h"

namespace gprt_host {

namespace {

constexpr float kAabbWidth = 2.0f;    // src/Renderer.h:95
constexpr float kAabbDistance = 2.0f; // src/Renderer.h:96

gprt_material make_material(const float albedo[4], float reflectance = 0.0f, float diffuse = 0.9f, float specular = 0.7f,
                            float power = 50.0f, float step_scale = 1.0f) { // set_attributes, src/Renderer.cpp:203-213
    gprt_material m;
    std::memset(&m, 0, sizeof m);
    for (int i = 0; i < 4; i++) m.albedo[i] = albedo[i];
    m.reflectance_coefficient = reflectance;
    m.diffuse_coefficient = diffuse;
    m.specular_coefficient = specular;
    m.specular_power = power;
    m.step_scale = step_scale;
    return m;
}

const float kGreen[4] = {0.1f, 1.0f, 0.5f, 1.0f};            // src/Renderer.cpp:218
const float kRed[4] = {1.0f, 0.5f, 0.5f, 1.0f};              // :219
const float kYellow[4] = {1.0f, 1.0f, 0.5f, 1.0f};           // :220
const float kChromium[4] = {0.549f, 0.556f, 0.554f, 1.0f};   // src/ConstantBuffers.h:135
const float kPlaneAlbedo[4] = {0.9f, 0.9f, 0.9f, 1.0f};      // src/Renderer.cpp:215

// The ten primitive templates of the reference, in instance_index order
// (src/Renderer.cpp:222-248 materials, :327-355 transforms, :505-535 AABB sizes).
struct Template {
    uint32_t kind, type;
    gprt_material material;
    float scale[3];
    bool rotates;
    float size[3];
};

void templates(Template t[10]) {
    t[0] = {GPRT_HG_AABB_ANALYTIC, GPRT_ANALYTIC_AABB, make_material(kRed), {1.0f, 1.5f, 1.0f}, false, {2, 3, 2}};
    t[1] = {GPRT_HG_AABB_ANALYTIC, GPRT_ANALYTIC_SPHERES, make_material(kChromium, 1), {1.5f, 1.5f, 1.5f}, true, {3, 3, 3}};
    t[2] = {GPRT_HG_AABB_VOLUMETRIC, GPRT_VOLUMETRIC_METABALLS, make_material(kChromium, 1), {1.5f, 1.5f, 1.5f}, true, {3, 3, 3}};
    t[3] = {GPRT_HG_AABB_SDF, GPRT_SDF_MINI_SPHERES, make_material(kGreen), {1, 1, 1}, false, {2, 2, 2}};
    t[4] = {GPRT_HG_AABB_SDF, GPRT_SDF_INTERSECTED_ROUND_CUBE, make_material(kGreen), {1, 1, 1}, false, {2, 2, 2}};
    t[5] = {GPRT_HG_AABB_SDF, GPRT_SDF_SQUARE_TORUS, make_material(kChromium, 1), {1.5f, 1.5f, 1.5f}, false, {3, 3, 3}};
    t[6] = {GPRT_HG_AABB_SDF, GPRT_SDF_TWISTED_TORUS, make_material(kYellow, 0, 1.0f, 0.7f, 50, 0.5f), {1, 1, 1}, true, {2, 2, 2}};
    t[7] = {GPRT_HG_AABB_SDF, GPRT_SDF_COG, make_material(kYellow, 0, 1.0f, 0.1f, 2), {1, 1, 1}, true, {2, 2, 2}};
    t[8] = {GPRT_HG_AABB_SDF, GPRT_SDF_CYLINDER, make_material(kRed), {1.0f, 1.5f, 1.0f}, false, {2, 3, 2}};
    t[9] = {GPRT_HG_AABB_SDF, GPRT_SDF_FRACTAL_PYRAMID, make_material(kGreen, 0, 1, 0.1f, 4, 0.8f), {3, 3, 3}, false, {6, 6, 6}};
}

// initialize_aabb, src/Renderer.cpp:498-504 (float arithmetic, same order).
gprt_aabb initialize_aabb(const float base[3], const float stride[3], float ox, float oy, float oz, const float size[3]) {
    gprt_aabb b;
    b.min_x = base[0] + ox * stride[0];
    b.min_y = base[1] + oy * stride[1];
    b.min_z = base[2] + oz * stride[2];
    b.max_x = base[0] + ox * stride[0] + size[0];
    b.max_y = base[1] + oy * stride[1] + size[1];
    b.max_z = base[2] + oz * stride[2] + size[2];
    return b;
}

PrimitiveSpec from_template(const Template& t, const gprt_aabb& box) {
    PrimitiveSpec p;
    p.hit_group_kind = t.kind;
    p.primitive_type = t.type;
    p.material = t.material;
    for (int i = 0; i < 3; i++) p.scale[i] = t.scale[i];
    p.rotates = t.rotates;
    p.aabb = box;
    return p;
}

void grid_base(int nx, int ny, int nz, float width, float distance, float base[3]) { // src/Renderer.cpp:491-495
    base[0] = -(nx * width + (nx - 1) * distance) / 2.0f;
    base[1] = -(ny * width + (ny - 1) * distance) / 2.0f;
    base[2] = -(nz * width + (nz - 1) * distance) / 2.0f;
}

void set_lights(SceneState& s) { // src/Renderer.cpp:271-286
    const float lp[4] = {0.0f, 18.0f, -20.0f, 0.0f}, la[4] = {0.25f, 0.25f, 0.25f, 1.0f}, ld[4] = {0.6f, 0.6f, 0.6f, 1.0f};
    for (int i = 0; i < 4; i++) {
        s.scene_cb.light_position[i] = lp[i];
        s.scene_cb.light_ambient_color[i] = la[i];
        s.scene_cb.light_diffuse_color[i] = ld[i];
    }
}

void set_plane(SceneState& s, uint32_t num_aabb_x, uint32_t num_aabb_z, float base_frac) { // src/Renderer.cpp:740-765
    const float fw[3] = {num_aabb_x * kAabbWidth + (num_aabb_x - 1) * kAabbDistance, 1 * kAabbWidth + 0 * kAabbDistance,
                         num_aabb_z * kAabbWidth + (num_aabb_z - 1) * kAabbDistance};
    for (int i = 0; i < 3; i++) s.plane_extent[i] = fw[i];
    s.plane_translate[0] = fw[0] * base_frac;
    s.plane_translate[1] = fw[1] * 0.0f;
    s.plane_translate[2] = fw[2] * base_frac;
    s.aabb_instance_translate[0] = 0.0f;
    s.aabb_instance_translate[1] = kAabbWidth / 2.0f; // src/Renderer.cpp:779
    s.aabb_instance_translate[2] = 0.0f;
}

void reference_primitives(SceneState& s) { // build_procedural_geometry_aabbs, src/Renderer.cpp:484-539
    Template t[10];
    templates(t);
    float base[3];
    grid_base(4, 1, 4, kAabbWidth, kAabbDistance, base);
    const float stride[3] = {kAabbWidth + kAabbDistance, kAabbWidth + kAabbDistance, kAabbWidth + kAabbDistance};
    const float off[10][3] = {{3, 0, 0}, {2.25f, 0, 0.75f}, {0, 0, 0}, {2, 0, 0}, {0, 0, 2}, {0.75f, -0.1f, 2.25f}, {0, 0, 1}, {1, 0, 0}, {0, 0, 3}, {2, 0, 2}};
    for (int i = 0; i < 10; i++)
        s.primitives.push_back(from_template(t[i], initialize_aabb(base, stride, off[i][0], off[i][1], off[i][2], t[i].size)));
}

} // namespace

void initialize_camera(SceneState& s, const double eye_in[3], const double at_in[3]) { // src/Renderer.cpp:253-268
    double eye[4] = {eye_in[0], eye_in[1], eye_in[2], 1.0};
    double at[4] = {at_in[0], at_in[1], at_in[2], 1.0};
    const double right[3] = {1.0, 0.0, 0.0};
    double direction[3] = {at[0] - eye[0], at[1] - eye[1], at[2] - eye[2]};
    vec3_normalize(direction);
    double up[3];
    vec3_cross(direction, right, up);
    vec3_normalize(up);
    const Mat4 rotate = mat_rotation_y(45.0 * M_PI / 180.0);
    vec3_transform(eye, rotate, s.eye);
    vec3_transform(up, rotate, s.up);
    for (int i = 0; i < 4; i++) s.at[i] = at[i];
    update_camera_matrices(s);
}

void update_camera_matrices(SceneState& s) { // src/Renderer.cpp:289-300
    for (int i = 0; i < 4; i++) s.scene_cb.camera_position[i] = (float)s.eye[i];
    const double aspect = (double)s.width / (double)s.height;
    const Mat4 view = mat_look_at_lh(s.eye, s.at, s.up);
    const Mat4 proj = mat_perspective_fov_lh((double)s.fov_angle_y_deg * M_PI / 180.0, aspect, s.z_near, s.z_far);
    mat_store(mat_inverse(mat_mul(view, proj)), s.scene_cb.projection_to_world);
}

void rotate_camera_y(SceneState& s, float degrees) { // src/Renderer.cpp:92-101
    const Mat4 rotate = mat_rotation_y((double)degrees * M_PI / 180.0);
    vec3_transform(s.eye, rotate, s.eye);
    vec3_transform(s.up, rotate, s.up);
    vec3_transform(s.at, rotate, s.at);
    update_camera_matrices(s);
}

void rotate_light_y(SceneState& s, float degrees) { // src/Renderer.cpp:104-110
    const Mat4 rotate = mat_rotation_y((double)degrees * M_PI / 180.0);
    const double l[3] = {s.scene_cb.light_position[0], s.scene_cb.light_position[1], s.scene_cb.light_position[2]};
    double o[4];
    vec3_transform(l, rotate, o);
    for (int i = 0; i < 4; i++) s.scene_cb.light_position[i] = (float)o[i];
}

void set_animation_time(SceneState& s, float animation_time) {
    s.animate_geometry_time = animation_time;
    s.scene_cb.elapsed_time = animation_time; // src/Renderer.cpp:119
}

void update_aabb_primitive_attributes(SceneState& s, float animation_time) { // src/Renderer.cpp:302-356
    s.animate_geometry_time = animation_time;
    s.scene_cb.elapsed_time = animation_time; // src/Renderer.cpp:119
    const Mat4 rotation = mat_rotation_y(-2.0 * (double)animation_time);
    const Mat4 identity = mat_identity();
    s.per_frame.resize(s.primitives.size());
    for (size_t i = 0; i < s.primitives.size(); i++) {
        const PrimitiveSpec& p = s.primitives[i];
        // v_translation = 0.5f * (min + max), float arithmetic as in the reference
        const float tx = 0.5f * (p.aabb.min_x + p.aabb.max_x), ty = 0.5f * (p.aabb.min_y + p.aabb.max_y), tz = 0.5f * (p.aabb.min_z + p.aabb.max_z);
        const Mat4 transform = mat_mul(mat_mul(mat_scaling(p.scale[0], p.scale[1], p.scale[2]), p.rotates ? rotation : identity), mat_translation(tx, ty, tz));
        mat_store(transform, s.per_frame[i].local_space_to_bottom_level_as);
        mat_store(mat_inverse(transform), s.per_frame[i].bottom_level_as_to_local_space);
    }
}

void build_geometry_and_tables(SceneState& s) {
    // build_plane_geometry, src/Renderer.cpp:541-565
    const uint16_t idx[6] = {3, 1, 0, 2, 1, 3};
    s.indices.assign(idx, idx + 6);
    const float pos[4][3] = {{0, 0, 0}, {1, 0, 0}, {1, 0, 1}, {0, 0, 1}};
    s.vertices.resize(4);
    for (int i = 0; i < 4; i++) {
        for (int k = 0; k < 3; k++) s.vertices[i].position[k] = pos[i][k];
        s.vertices[i].normal[0] = 0; s.vertices[i].normal[1] = 1; s.vertices[i].normal[2] = 0;
    }
    // AABBs
    s.aabbs.resize(s.primitives.size());
    s.anim.resize(s.primitives.size());
    for (size_t i = 0; i < s.primitives.size(); i++) {
        s.aabbs[i] = s.primitives[i].aabb;
        for (int k = 0; k < 3; k++) s.anim[i].scale[k] = s.primitives[i].scale[k];
        s.anim[i].rotates = s.primitives[i].rotates ? 1u : 0u;
    }

    // build_bottom_level_as_instance_descs, src/Renderer.cpp:732-786
    s.instances.assign(2, gprt_instance_desc());
    std::memset(s.instances.data(), 0, 2 * sizeof(gprt_instance_desc));
    {
        gprt_instance_desc& d = s.instances[0];
        const Mat4 m = mat_mul(mat_scaling(s.plane_extent[0], s.plane_extent[1], s.plane_extent[2]),
                               mat_translation(s.plane_translate[0], s.plane_translate[1], s.plane_translate[2]));
        mat_store_3x4(m, d.transform);
        d.instance_id_and_mask = (1u << 24);          // InstanceMask = 1
        d.hit_group_contrib_and_flags = 0;            // InstanceContributionToHitGroupIndex = 0
        d.acceleration_structure = GPRT_BLAS_TRIANGLES;
    }
    {
        gprt_instance_desc& d = s.instances[1];
        mat_store_3x4(mat_translation(s.aabb_instance_translate[0], s.aabb_instance_translate[1], s.aabb_instance_translate[2]), d.transform);
        d.instance_id_and_mask = (1u << 24);
        d.hit_group_contrib_and_flags = 2;            // BottomLevelASType::AABB * RayType::Count
        d.acceleration_structure = GPRT_BLAS_AABBS;
    }

    // build_shader_tables (hit-group table), src/Renderer.cpp:936-984
    const size_t n = s.primitives.size();
    s.hit_groups.assign(2 + 2 * n, gprt_hit_group_record());
    std::memset(s.hit_groups.data(), 0, s.hit_groups.size() * sizeof(gprt_hit_group_record));
    for (uint32_t r = 0; r < 2; r++) {
        gprt_hit_group_record& rec = s.hit_groups[r];
        rec.shader_id[0] = GPRT_HG_TRIANGLE;
        rec.shader_id[1] = r;
        rec.material_cb = s.plane_material;
    }
    for (size_t i = 0; i < n; i++) {
        for (uint32_t r = 0; r < 2; r++) {
            gprt_hit_group_record& rec = s.hit_groups[2 + 2 * i + r];
            rec.shader_id[0] = s.primitives[i].hit_group_kind;
            rec.shader_id[1] = r;
            rec.material_cb = s.primitives[i].material;
            rec.aabb_cb.instance_index = (uint32_t)i;
            rec.aabb_cb.primitive_type = s.primitives[i].primitive_type;
        }
    }
    if (!s.extra_instances) return;

    // Two-level instancing (not in the reference, which has one instance per BLAS): two more instances of the AABB
    // BLAS -- one scaled and rotated that shares the first one's hit groups, one whose
    // InstanceContributionToHitGroupIndex selects a second block of records (other materials, and the two
    // 2x2x2 SDF primitives MiniSpheres / IntersectedRoundCube swapped, so a primitive's intersection shader
    // depends on the instance) -- and a second instance of the quad as a raised mirror panel with its own record.
    const uint32_t block_b = (uint32_t)(2 + 2 * n), panel = (uint32_t)(2 + 4 * n);
    s.hit_groups.resize(2 + 4 * n + 2);
    for (size_t i = 0; i < n; i++) {
        for (uint32_t r = 0; r < 2; r++) {
            gprt_hit_group_record rec = s.hit_groups[2 + 2 * i + r];
            const float a0 = rec.material_cb.albedo[0];
            rec.material_cb.albedo[0] = rec.material_cb.albedo[2];
            rec.material_cb.albedo[2] = a0;
            rec.material_cb.reflectance_coefficient = rec.material_cb.reflectance_coefficient > 0.5f ? 0.0f : 0.3f;
            if (rec.shader_id[0] == GPRT_HG_AABB_SDF && rec.aabb_cb.primitive_type == GPRT_SDF_MINI_SPHERES) rec.aabb_cb.primitive_type = GPRT_SDF_INTERSECTED_ROUND_CUBE;
            else if (rec.shader_id[0] == GPRT_HG_AABB_SDF && rec.aabb_cb.primitive_type == GPRT_SDF_INTERSECTED_ROUND_CUBE) rec.aabb_cb.primitive_type = GPRT_SDF_MINI_SPHERES;
            s.hit_groups[block_b + 2 * i + r] = rec;
        }
    }
    const float red[4] = {1.0f, 0.45f, 0.4f, 1.0f};
    for (uint32_t r = 0; r < 2; r++) {
        gprt_hit_group_record rec = s.hit_groups[r];
        rec.material_cb = make_material(red, 0.6f, 0.9f, 0.7f, 50, 1);
        s.hit_groups[panel + r] = rec;
    }
    const float* tr = s.aabb_instance_translate;
    gprt_instance_desc d;
    std::memset(&d, 0, sizeof d);
    d.instance_id_and_mask = (1u << 24) | 1u;
    d.hit_group_contrib_and_flags = 2;
    d.acceleration_structure = GPRT_BLAS_AABBS;
    mat_store_3x4(mat_mul(mat_mul(mat_scaling(0.7, 0.7, 0.7), mat_rotation_y(0.6)), mat_translation(tr[0] + 17.0, tr[1] + 0.0, tr[2] + 5.0)), d.transform);
    s.instances.push_back(d);
    d.instance_id_and_mask = (1u << 24) | 2u;
    d.hit_group_contrib_and_flags = block_b;
    mat_store_3x4(mat_mul(mat_rotation_y(-0.9), mat_translation(tr[0] - 14.0, tr[1] + 0.0, tr[2] + 9.0)), d.transform);
    s.instances.push_back(d);
    d.instance_id_and_mask = (1u << 24) | 3u;
    d.hit_group_contrib_and_flags = panel;
    d.acceleration_structure = GPRT_BLAS_TRIANGLES;
    mat_store_3x4(mat_mul(mat_scaling(8.0, 1.0, 8.0), mat_translation(-4.0, 4.0, 14.0)), d.transform);
    s.instances.push_back(d);
}

bool build_scene(SceneState& s, int scene_id, uint32_t width, uint32_t height, uint32_t grid_n) {
    s = SceneState();
    gprt_default_config(&s.config);
    std::memset(&s.scene_cb, 0, sizeof s.scene_cb);
    s.plane_material = make_material(kPlaneAlbedo, 0.25f, 1, 0.4f, 50, 1); // src/Renderer.cpp:215
    set_lights(s);
    set_plane(s, 700, 700, -0.35f);
    double eye[3] = {0.0, 5.3, -17.0}, at[3] = {0.0, 0.0, 0.0}; // src/Renderer.cpp:256-257
    float time0 = 0.0f;
    Template t[10];
    templates(t);

    switch (scene_id) {
    case kSceneReference:
        s.width = width ? width : 1280; s.height = height ? height : 720;
        reference_primitives(s);
        break;
    case kConfig2:
        s.width = width ? width : 1920; s.height = height ? height : 1080;
        reference_primitives(s);
        break;
    case kInstanced:
        s.width = width ? width : 1280; s.height = height ? height : 720;
        reference_primitives(s);
        s.extra_instances = true;
        eye[1] = 9.0; eye[2] = -36.0;
        break;
    case kConfig1: {
        // one analytic sphere-cluster AABB + the ground quad, primary + shadow rays only:
        // depth 2 (under the reference's depth rule depth 1 traces no shadow rays) and no
        // reflective material anywhere.
        s.width = width ? width : 640; s.height = height ? height : 480;
        s.config.max_recursion_depth = 2;
        s.plane_material.reflectance_coefficient = 0.0f;
        float base[3];
        grid_base(4, 1, 4, kAabbWidth, kAabbDistance, base);
        const float stride[3] = {4, 4, 4};
        Template sp = t[1];
        sp.material.reflectance_coefficient = 0.0f;
        s.primitives.push_back(from_template(sp, initialize_aabb(base, stride, 1.5f, 0, 1.5f, sp.size)));
        break;
    }
    case kConfig3:
    case kConfig3Long: {
        // SDF-heavy: 4x1x4 cells, FractalPyramid in every second cell, the other six SDF types
        // cycling through the rest; reference per-type scale, material and AABB size.
        s.width = width ? width : 3840; s.height = height ? height : 2160;
        if (scene_id == kConfig3Long) { s.config.sdf_max_steps = 2048; s.config.metaball_max_steps = 512; }
        const float cell = 8.0f;
        const float base[3] = {-2.0f * cell + 1.0f, -1.0f, -2.0f * cell + 1.0f};
        const float stride[3] = {cell, cell, cell};
        const int others[6] = {3, 4, 5, 6, 7, 8};
        int k = 0;
        for (int iz = 0; iz < 4; iz++)
            for (int ix = 0; ix < 4; ix++) {
                const bool pyramid = ((ix + iz) & 1) == 0;
                const Template& tt = pyramid ? t[9] : t[others[(k++) % 6]];
                s.primitives.push_back(from_template(tt, initialize_aabb(base, stride, (float)ix, 0, (float)iz, tt.size)));
            }
        eye[1] = 12.0; eye[2] = -38.0;
        break;
    }
    case kConfig4: {
        // 3x1x3 metaball AABBs, METABALLS_COUNT == 5 variant of the reference.
        s.width = width ? width : 1920; s.height = height ? height : 1080;
        s.config.metaballs_count = 5;
        float base[3];
        grid_base(3, 1, 3, kAabbWidth, kAabbDistance, base);
        const float stride[3] = {4, 4, 4};
        for (int iz = 0; iz < 3; iz++)
            for (int ix = 0; ix < 3; ix++)
                s.primitives.push_back(from_template(t[2], initialize_aabb(base, stride, (float)ix, 0, (float)iz, t[2].size)));
        break;
    }
    case kConfig5: {
        // n x 1 x n instanced grid (n = 1000 -> 1M primitives), all ten reference primitive
        // types cycling, unit scale so every primitive fits its 2x2x2 cell.
        s.width = width ? width : 7680; s.height = height ? height : 4320;
        const int n = grid_n ? (int)grid_n : 1000;
        float base[3];
        grid_base(n, 1, n, kAabbWidth, kAabbDistance, base);
        const float stride[3] = {4, 4, 4};
        const float size[3] = {2, 2, 2};
        s.primitives.reserve((size_t)n * n);
        for (int iz = 0; iz < n; iz++)
            for (int ix = 0; ix < n; ix++) {
                Template tt = t[(ix + 3 * iz) % 10];
                tt.scale[0] = tt.scale[1] = tt.scale[2] = 1.0f;
                s.primitives.push_back(from_template(tt, initialize_aabb(base, stride, (float)ix, 0, (float)iz, size)));
            }
        set_plane(s, (uint32_t)(n * 1.4) + 1, (uint32_t)(n * 1.4) + 1, -0.5f);
        eye[1] = 25.0; eye[2] = -50.0; // fog (exp(-2e-6 t^3)) hides everything beyond t ~ 125
        time0 = 1.0f;
        break;
    }
    default:
        return false;
    }
    initialize_camera(s, eye, at);
    build_geometry_and_tables(s);
    update_aabb_primitive_attributes(s, time0);
    return true;
}

} // namespace gprt_host
