// megakernel.cuh -- the comparison point SURVEY.md 7.3 asks to measure beside the wavefront form: ONE kernel, one thread
// per pixel, the pixel's whole ray tree in registers, no queues and no HBM traffic between the stages
// (GPRT_FLAG_MEGAKERNEL; included by kernels.cu, same anonymous namespace).
//
// It is the reference's own shape -- MyRaygenShader -> TraceRay -> intersection / closest-hit shaders recursing to
// MAX_RAY_RECURSION_DEPTH (src/Raytracing.hlsl:154-253) -- with the recursion unrolled into a level loop and the colour
// folded back to front, and the same arithmetic as the wavefront kernels (prims.cuh / shade.cuh), so the two produce the
// same image bit for bit (tests/test_gpu_parity.py).  Every AABB is tested by every ray (the scenes it is meant for have a
// dozen primitives; GPRT_E_UNSUPPORTED beyond kMegaMaxPrims or with several TLAS instances of a kind).  What it costs is
// what the wavefront form exists to avoid: the 32 pixels of a warp march different primitive types one after the other, a
// warp is as slow as its slowest pixel, and one pixel's three levels are a single dependent chain (measured: DESIGN.md 4).
#pragma once

namespace gprt {
namespace {

constexpr uint32_t kMegaMaxPrims = 64;

// MyIntersectionShader_* for primitive `prim` on the object-space ray (oo, dd) (src/Raytracing.hlsl:277-341): transforms to
// the primitive's local space, runs its test, returns the hit distance and the world normal.
template <int NB>
__device__ __forceinline__ bool mega_intersect(const SceneDev& sc, const PassParams& pp, uint32_t prim, V3 oo, V3 dd, float tcur, float& thit, V3& nw,
                                               uint32_t& work) {
    const PrimRecord pr = sc.prims[prim];
    const float* pf = sc.per_frame + 32 * (size_t)pr.instance_index;
    LocalRay lr;
    lr.o = mul_point_rowvec(oo, pf + 16);
    lr.d = mul_vector_rowvec(dd, pf + 16);
    lr.tmin = 0.0f;
    lr.tcur = tcur;
    V3 n = v3s(0.0f);
    bool hit = false;
    switch (pr.type_id) {
    case 0: hit = run_test<0, NB>(lr, pr.mat, pp, thit, n, work); break;
    case 1: hit = run_test<1, NB>(lr, pr.mat, pp, thit, n, work); break;
    case 2: hit = run_test<2, NB>(lr, pr.mat, pp, thit, n, work); break;
    case 3: hit = run_test<3, NB>(lr, pr.mat, pp, thit, n, work); break;
    case 4: hit = run_test<4, NB>(lr, pr.mat, pp, thit, n, work); break;
    case 5: hit = run_test<5, NB>(lr, pr.mat, pp, thit, n, work); break;
    case 6: hit = run_test<6, NB>(lr, pr.mat, pp, thit, n, work); break;
    case 7: hit = run_test<7, NB>(lr, pr.mat, pp, thit, n, work); break;
    case 8: hit = run_test<8, NB>(lr, pr.mat, pp, thit, n, work); break;
    default: hit = run_test<9, NB>(lr, pr.mat, pp, thit, n, work); break;
    }
    if (!(hit && thit >= lr.tmin && thit <= lr.tcur)) return false;
    const V3 nb = mul_vector_rowvec(n, pf);
    nw = normalize(xform_vector_3x4(sc.aabb_o2w, nb));
    return true;
}

// TraceRay for one ray: the plane triangles, then every AABB whose slab test passes.  any_hit: a shadow ray
// (ACCEPT_FIRST_HIT_AND_END_SEARCH); otherwise the closest hit, ties to the lowest primitive like the wavefront's key.
template <int NB>
__device__ __forceinline__ unsigned long long mega_trace(const SceneDev& sc, const PassParams& pp, V3 o, V3 d, bool any_hit, V3& n_out, uint32_t& work) {
    float tcur = kRayTMax;
    unsigned long long key = kKeyMiss;
    plane_triangles<false>(sc, o, d, any_hit, tcur, key);
    if (any_hit && key != kKeyMiss) return key;
    if (!sc.has_aabb_instance) return key;
    const V3 oo = xform_point_3x4(sc.aabb_w2o, o), dd = xform_vector_3x4(sc.aabb_w2o, d);
    const V3 inv_d = v3(1.0f / dd.x, 1.0f / dd.y, 1.0f / dd.z);
    for (uint32_t i = 0; i < sc.n_prims; i++) {
        const float* b = sc.aabbs + 6 * i;
        if (!slab_test(oo, inv_d, b, b + 3, 0.0f, tcur)) continue;
        float thit = 0.0f;
        V3 nw = v3s(0.0f);
        if (!mega_intersect<NB>(sc, pp, i, oo, dd, tcur, thit, nw, work)) continue;
        const unsigned long long k = pack_key(thit, i);
        if (k < key) { key = k; n_out = nw; }
        if (any_hit) return key;
    }
    return key;
}

template <int NB>
__global__ void __launch_bounds__(128) k_megakernel(const __grid_constant__ SceneDev sc, const __grid_constant__ PassParams pp,
                                                    const __grid_constant__ WaveBuffers wb, const __grid_constant__ OutputBuffers ob) {
    uint32_t n_radiance[kMaxDepth] = {}, n_shadow[kMaxDepth] = {};
    uint32_t work = 0;
    for (uint32_t pixel = blockIdx.x * blockDim.x + threadIdx.x; pixel < pp.n_pixels; pixel += gridDim.x * blockDim.x) {
        const uint32_t tx = pixel % pp.tile_w, ty = pixel / pp.tile_w;
        const uint32_t px = pp.tile_x + tx, py = pass_global_row(pp, ty);
        // the levels of this pixel's ray tree (a mirror bounce per level), front to back
        V3 lo[kMaxDepth], ld[kMaxDepth], ln[kMaxDepth];
        float lt[kMaxDepth];
        uint32_t lcode[kMaxDepth];
        bool lshadow[kMaxDepth];
        V3 o, d;
        generate_camera_ray(px, py, pp.fc, o, d);
        uint32_t levels = 0;
        for (uint32_t level = 1; level <= pp.max_depth; level++) { // TraceRadianceRay (src/Raytracing.hlsl:87-112)
            V3 n = v3s(0.0f);
            const unsigned long long key = mega_trace<NB>(sc, pp, o, d, false, n, work);
            n_radiance[level - 1]++;
            const uint32_t k = levels++;
            lo[k] = o; ld[k] = d; lshadow[k] = false;
            if (key == kKeyMiss) { lcode[k] = kCodeMiss; break; }
            const uint32_t code = (uint32_t)(key & 0xFFFFFFFFull);
            const float t = __uint_as_float((uint32_t)(key >> 32));
            float refl;
            if (code & kCodePlane) {
                const float* tn = sc.tri_nrm + 3 * (code & ~kCodePlane);
                n = v3(tn[0], tn[1], tn[2]);
                refl = sc.plane_mat.reflectance_coefficient;
            } else {
                refl = sc.prims[code].mat.reflectance_coefficient;
            }
            lcode[k] = code; lt[k] = t; ln[k] = n;
            if (level >= pp.max_depth) break; // the closest-hit shaders' own TraceRay calls return at the depth limit (:89,:117)
            const V3 hit_pos = o + t * d;
            const V3 lp = v3(pp.fc.light_position[0], pp.fc.light_position[1], pp.fc.light_position[2]);
            V3 ns = v3s(0.0f);
            lshadow[k] = mega_trace<NB>(sc, pp, hit_pos, normalize(lp - hit_pos), true, ns, work) != kKeyMiss; // TraceShadowRayAndReportIfHit
            n_shadow[level - 1]++;
            if (!(refl > 0.001f)) break;
            d = reflect(d, n);
            o = hit_pos;
        }
        V4 color = v4s(0.0f);
        for (int k = (int)levels - 1; k >= 0; k--) {
            if (lcode[k] == kCodeMiss) { color = background_color(); continue; }
            const bool is_tri = (lcode[k] & kCodePlane) != 0u;
            const Material mat = is_tri ? sc.plane_mat : sc.prims[lcode[k]].mat;
            color = closest_hit_color(pp.fc, lo[k], ld[k], lt[k], ln[k], mat, lshadow[k], color, is_tri, px, py);
        }
        const uint32_t ox = ob.out_x0 + tx, oy = ob.out_y0 + ty;
        *reinterpret_cast<uint32_t*>(ob.rgba8 + (size_t)oy * ob.pitch + (size_t)ox * 4u) = pack_unorm8(color);
        const size_t oi = (size_t)oy * ob.out_w + ox;
        if (ob.rgba_f32) ob.rgba_f32[oi] = make_float4(color.x, color.y, color.z, color.w);
        if (ob.hit_ids) {
            const uint32_t code = lcode[0];
            ob.hit_ids[oi] = (code == kCodeMiss) ? 0xFFFFFFFFu : (code & kCodePlane) ? (0x00000000u | (code & ~kCodePlane)) : (0x10000000u | code);
        }
    }
    // ray statistics: wave k of the wavefront form = level-k radiance rays + level-(k-1) shadow rays
    for (uint32_t l = 0; l < pp.max_depth; l++) {
        const uint32_t r = __reduce_add_sync(kFull, n_radiance[l]), s = __reduce_add_sync(kFull, n_shadow[l]);
        if ((threadIdx.x & 31) == 0) {
            if (r) atomicAdd(&wb.ctrl[l + 1].n_radiance, r);
            if (s) atomicAdd(&wb.ctrl[l + 2].n_shadow, s);
        }
    }
    (void)work;
}

} // namespace
} // namespace gprt
