// shade.cuh -- camera rays and surface shading (device-side counterparts of the raygen,
// closest-hit and miss shaders).
//   GenerateCameraRay            src/RaytracingShaderHelper.hlsli:103-120
//   CalculatePhongLighting etc.  src/Raytracing.hlsl:35-80
//   FresnelReflectanceSchlick    src/RaytracingShaderHelper.hlsli:178-182
//   AnalyticalCheckersTexture    src/RaytracingShaderHelper.hlsli:148-175,
//   CheckersTextureBoxFilter     src/SignedDistancePrimitives.hlsli:324-334
//   MyClosestHitShader_Triangle / _AABB, MyMissShader   src/Raytracing.hlsl:172-264
#pragma once
#include "vecmath.cuh"

namespace gprt {

struct FrameConstants {           // SceneConstantBuffer + dispatch dimensions
    float projection_to_world[16];
    float camera_position[4];
    float light_position[4];
    float light_ambient_color[4];
    float light_diffuse_color[4];
    float elapsed_time;
    uint32_t width, height;       // DispatchRaysDimensions().xy
};

struct Material {                 // PrimitiveConstantBuffer
    float albedo[4];
    float reflectance_coefficient, diffuse_coefficient, specular_coefficient, specular_power, step_scale;
};

GPRT_HD void generate_camera_ray(uint32_t px, uint32_t py, const FrameConstants& fc, V3& origin, V3& direction) {
    const float x = (float)px + 0.5f, y = (float)py + 0.5f;
    float sx = x / (float)fc.width * 2.0f - 1.0f;
    float sy = y / (float)fc.height * 2.0f - 1.0f;
    sy = -sy;
    const float* m = fc.projection_to_world;
    // mul(float4(screenPos, 0, 1), projectionToWorld)
    const float wx = sx * m[0] + sy * m[4] + 0.0f * m[8] + 1.0f * m[12];
    const float wy = sx * m[1] + sy * m[5] + 0.0f * m[9] + 1.0f * m[13];
    const float wz = sx * m[2] + sy * m[6] + 0.0f * m[10] + 1.0f * m[14];
    const float ww = sx * m[3] + sy * m[7] + 0.0f * m[11] + 1.0f * m[15];
    const V3 world = v3(wx / ww, wy / ww, wz / ww);
    origin = v3(fc.camera_position[0], fc.camera_position[1], fc.camera_position[2]);
    direction = normalize(world - origin);
}

GPRT_HD V3 fresnel_schlick(V3 I, V3 N, V3 f0) {
    const float cosi = saturate(dot(-I, N));
    const float p = hlsl_pow(1.0f - cosi, 5.0f);
    return f0 + (v3s(1.0f) - f0) * v3s(p);
}

GPRT_HD float checkers_box_filter(V2 uv, V2 dpdx, V2 dpdy, uint32_t ratio) {
    const float fr = (float)ratio;
    const V2 w = v2(fmaxf(fabsf(dpdx.x), fabsf(dpdy.x)), fmaxf(fabsf(dpdx.y), fabsf(dpdy.y)));
    const V2 a = uv + 0.5f * w;
    const V2 b = uv - 0.5f * w;
    const float ix = (floorf(a.x) + fminf(frac(a.x) * fr, 1.0f) - floorf(b.x) - fminf(frac(b.x) * fr, 1.0f)) / (fr * w.x);
    const float iy = (floorf(a.y) + fminf(frac(a.y) * fr, 1.0f) - floorf(b.y) - fminf(frac(b.y) * fr, 1.0f)) / (fr * w.y);
    return (1.0f - ix) * (1.0f - iy);
}

// AnalyticalCheckersTexture: ray differentials of the PRIMARY pixel (DispatchRaysIndex), also
// for reflected hits.
GPRT_HD float analytical_checkers(V3 hit_position, V3 n, uint32_t px, uint32_t py, const FrameConstants& fc) {
    const V2 uv = v2(hit_position.x, hit_position.z);
    V3 ox, dx, oy, dy;
    generate_camera_ray(px + 1u, py, fc, ox, dx);
    generate_camera_ray(px, py + 1u, fc, oy, dy);
    const V3 ddx_pos = ox - dx * v3s(dot(ox - hit_position, n)) / v3s(dot(dx, n));
    const V3 ddy_pos = oy - dy * v3s(dot(oy - hit_position, n)) / v3s(dot(dy, n));
    const V2 ddx_uv = v2(ddx_pos.x, ddx_pos.z) - uv;
    const V2 ddy_uv = v2(ddy_pos.x, ddy_pos.z) - uv;
    return checkers_box_filter(uv, ddx_uv, ddy_uv, 50u);
}

GPRT_HD V4 phong_lighting(const FrameConstants& fc, V3 hit_position, V3 ray_direction, const Material& mat, V3 normal, bool in_shadow) {
    const V4 albedo = v4(mat.albedo[0], mat.albedo[1], mat.albedo[2], mat.albedo[3]);
    const V3 light_position = v3(fc.light_position[0], fc.light_position[1], fc.light_position[2]);
    const float shadow_factor = in_shadow ? 0.35f : 1.0f; // IN_SHADOW_RADIANCE
    const V3 incident = normalize(hit_position - light_position);

    const V4 light_diffuse = v4(fc.light_diffuse_color[0], fc.light_diffuse_color[1], fc.light_diffuse_color[2], fc.light_diffuse_color[3]);
    const float kd = saturate(dot(-incident, normal));
    const V4 diffuse = ((shadow_factor * mat.diffuse_coefficient * kd) * light_diffuse) * albedo;

    V4 specular = v4s(0.0f);
    if (!in_shadow) {
        const V3 reflected = normalize(reflect(incident, normal));
        const float ks = hlsl_pow(saturate(dot(reflected, normalize(-ray_direction))), mat.specular_power);
        specular = (mat.specular_coefficient * v4s(ks)) * v4s(1.0f);
    }

    const V4 amb = v4(fc.light_ambient_color[0], fc.light_ambient_color[1], fc.light_ambient_color[2], fc.light_ambient_color[3]);
    const V4 amb_min = amb - v4s(0.1f);
    const float a = 1.0f - saturate(dot(normal, v3(0.0f, -1.0f, 0.0f)));
    const V4 ambient = albedo * lerp(amb_min, amb, a);
    return (ambient + diffuse) + specular;
}

GPRT_HD V4 background_color() { return v4(0.8f, 0.9f, 1.0f, 1.0f); }

// The body shared by MyClosestHitShader_Triangle / _AABB after their TraceRay calls returned.
//   is_triangle: multiply by the checker texture (plane only)
//   reflection_color: what TraceRadianceRay returned (zero when not traced)
GPRT_HD V4 closest_hit_color(const FrameConstants& fc, V3 ray_origin, V3 ray_direction, float t, V3 normal, const Material& mat,
                             bool in_shadow, V4 reflection_color, bool is_triangle, uint32_t px, uint32_t py) {
    const V3 hit_position = ray_origin + t * ray_direction;
    float checkers = 1.0f;
    if (is_triangle) checkers = analytical_checkers(hit_position, normal, px, py, fc);
    V4 reflected = v4s(0.0f);
    if (mat.reflectance_coefficient > 0.001f) {
        const V3 fr = fresnel_schlick(ray_direction, normal, v3(mat.albedo[0], mat.albedo[1], mat.albedo[2]));
        reflected = (mat.reflectance_coefficient * v4(fr.x, fr.y, fr.z, 1.0f)) * reflection_color;
    }
    const V4 phong = phong_lighting(fc, hit_position, ray_direction, mat, normal, in_shadow);
    V4 color = phong + reflected;
    if (is_triangle) color = checkers * color;
    const float fog = 1.0f - hlsl_exp(-0.000002f * t * t * t);
    return lerp(color, background_color(), fog);
}

// float4 -> R8G8B8A8_UNORM store of the output UAV (NaN -> 0, clamp, round to nearest)
GPRT_HD uint32_t pack_unorm8(V4 c) {
    float v[4] = {c.x, c.y, c.z, c.w};
    uint32_t out = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        float f = v[i];
        if (!(f == f)) f = 0.0f;
        f = f < 0.0f ? 0.0f : (f > 1.0f ? 1.0f : f);
        out |= ((uint32_t)floorf(f * 255.0f + 0.5f)) << (8 * i);
    }
    return out;
}

} // namespace gprt
