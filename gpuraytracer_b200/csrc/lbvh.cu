// lbvh.cu -- hand-written LBVH builder over the procedural AABBs.
//
// Replaces what Renderer::build_acceleration_structures (src/Renderer.cpp:788-835) asks the
// D3D12 driver for (BuildRaytracingAccelerationStructure on the AABB BLAS, :660,:722).  B200 has
// no RT cores, so the structure is ours:
//   1. scene bounds            block reduction + ordered-int atomics
//   2. 30-bit Morton codes     of the AABB centroids
//   3. LSD radix sort          4 passes x 8 bits, warp-private digit counters, stable
//   4. Karras 2012             one thread per internal node, (code, index) keys so duplicates split
//   5. bottom-up refit         second-arrival-continues, child boxes stored in the parent
// Node layout (64 B, 4 x float4): child0 min.xyz max.xyz | child1 min.xyz max.xyz | ref0 ref1 - -
// with ref >= 0 an internal node, ref < 0 a leaf (~primitive index).  One 64-byte fetch per visit
// yields both child boxes.
#include <cuda_runtime.h>
#include <stdint.h>

#include "lbvh.h"

namespace gprt {
namespace {

constexpr int kBlock = 256;
constexpr int kWarpsPerBlock = kBlock / 32;
constexpr int kKeysPerWarp = 32 * 16; // each warp owns a contiguous run of 512 keys
constexpr unsigned kFull = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t float_to_ordered(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

// bounds[0..2] = min (ordered), bounds[3..5] = max (ordered) of the centroids
__global__ void k_bounds(const float* __restrict__ aabbs, uint32_t n, uint32_t* bounds) {
    float lo[3] = {__builtin_inff(), __builtin_inff(), __builtin_inff()}, hi[3] = {-__builtin_inff(), -__builtin_inff(), -__builtin_inff()};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float* b = aabbs + 6 * (size_t)i;
        for (int k = 0; k < 3; k++) {
            const float c = 0.5f * (b[k] + b[k + 3]);
            lo[k] = fminf(lo[k], c);
            hi[k] = fmaxf(hi[k], c);
        }
    }
    for (int k = 0; k < 3; k++) {
        for (int off = 16; off; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(kFull, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(kFull, hi[k], off));
        }
    }
    if ((threadIdx.x & 31) == 0) {
        for (int k = 0; k < 3; k++) {
            atomicMin(&bounds[k], float_to_ordered(lo[k]));
            atomicMax(&bounds[3 + k], float_to_ordered(hi[k]));
        }
    }
}

__device__ __forceinline__ uint32_t expand_bits10(uint32_t v) { // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void k_morton(const float* __restrict__ aabbs, uint32_t n, const uint32_t* __restrict__ bounds, uint32_t* keys, uint32_t* vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* b = aabbs + 6 * (size_t)i;
    uint32_t q[3];
    for (int k = 0; k < 3; k++) {
        const float lo = ordered_to_float(bounds[k]), hi = ordered_to_float(bounds[3 + k]);
        const float ext = hi - lo;
        const float c = 0.5f * (b[k] + b[k + 3]);
        float u = ext > 0.0f ? (c - lo) / ext : 0.0f;
        u = fminf(fmaxf(u * 1024.0f, 0.0f), 1023.0f);
        q[k] = (uint32_t)u;
    }
    keys[i] = (expand_bits10(q[0]) << 2) | (expand_bits10(q[1]) << 1) | expand_bits10(q[2]);
    vals[i] = i;
}

// ---- radix sort: one pass = histogram, scan, scatter -----------------------------------------
// Warp w (global id) owns keys [w*512, w*512+512).  hist is digit-major: hist[d * n_warps + w].
__global__ void k_radix_hist(const uint32_t* __restrict__ keys, uint32_t n, int shift, uint32_t* hist, uint32_t n_warps) {
    __shared__ uint32_t s_cnt[kWarpsPerBlock][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t gw = blockIdx.x * kWarpsPerBlock + warp;
    for (int d = lane; d < 256; d += 32) s_cnt[warp][d] = 0;
    __syncwarp();
    if (gw < n_warps) {
        const uint32_t begin = gw * kKeysPerWarp;
        for (int it = 0; it < kKeysPerWarp / 32; it++) {
            const uint32_t i = begin + it * 32 + lane;
            const bool valid = i < n;
            const uint32_t d = valid ? ((keys[i] >> shift) & 255u) : 256u + lane; // invalid lanes never match
            const unsigned peers = __match_any_sync(kFull, d);
            if (valid && (__ffs(peers) - 1) == lane) s_cnt[warp][d] += __popc(peers);
            __syncwarp();
        }
        for (int d = lane; d < 256; d += 32) hist[(size_t)d * n_warps + gw] = s_cnt[warp][d];
    }
}

// exclusive scan of `m` counters by one block
__global__ void k_scan_exclusive(uint32_t* data, uint32_t m) {
    __shared__ uint32_t s_part[1024];
    const uint32_t per = (m + blockDim.x - 1) / blockDim.x;
    const uint32_t begin = threadIdx.x * per, end = min(begin + per, m);
    uint32_t sum = 0;
    for (uint32_t i = begin; i < end; i++) sum += data[i];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    for (uint32_t off = 1; off < blockDim.x; off <<= 1) { // Hillis-Steele inclusive
        uint32_t v = 0;
        if (threadIdx.x >= off) v = s_part[threadIdx.x - off];
        __syncthreads();
        s_part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = (threadIdx.x == 0) ? 0u : s_part[threadIdx.x - 1];
    for (uint32_t i = begin; i < end; i++) {
        const uint32_t v = data[i];
        data[i] = run;
        run += v;
    }
}

__global__ void k_radix_scatter(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, uint32_t* keys_out,
                                uint32_t* vals_out, uint32_t n, int shift, const uint32_t* __restrict__ hist, uint32_t n_warps) {
    __shared__ uint32_t s_off[kWarpsPerBlock][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t gw = blockIdx.x * kWarpsPerBlock + warp;
    if (gw >= n_warps) return;
    for (int d = lane; d < 256; d += 32) s_off[warp][d] = hist[(size_t)d * n_warps + gw];
    __syncwarp();
    const uint32_t begin = gw * kKeysPerWarp;
    for (int it = 0; it < kKeysPerWarp / 32; it++) {
        const uint32_t i = begin + it * 32 + lane;
        const bool valid = i < n;
        uint32_t key = 0, val = 0;
        if (valid) { key = keys_in[i]; val = vals_in[i]; }
        const uint32_t d = valid ? ((key >> shift) & 255u) : 256u + lane;
        const unsigned peers = __match_any_sync(kFull, d);
        uint32_t base = 0;
        if (valid) base = s_off[warp][d];
        __syncwarp();
        if (valid && (__ffs(peers) - 1) == lane) s_off[warp][d] = base + __popc(peers);
        __syncwarp();
        if (valid) {
            const uint32_t pos = base + __popc(peers & ((1u << lane) - 1u)); // stable inside the warp
            keys_out[pos] = key;
            vals_out[pos] = val;
        }
    }
}

// ---- Karras 2012 ------------------------------------------------------------------------------
__device__ __forceinline__ int delta(const uint32_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint32_t a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clz(a ^ b);
}

// children[2*i], children[2*i+1]: refs (>= 0 internal, < 0 leaf: ~sorted position); parents for refit
__global__ void k_karras(const uint32_t* __restrict__ keys, int n, int32_t* children, int32_t* parent_internal, int32_t* parent_leaf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int32_t left = (lo == gamma) ? ~gamma : gamma;
    const int32_t right = (hi == gamma + 1) ? ~(gamma + 1) : (gamma + 1);
    children[2 * i] = left;
    children[2 * i + 1] = right;
    if (left < 0) parent_leaf[~left] = i; else parent_internal[left] = i;
    if (right < 0) parent_leaf[~right] = i; else parent_internal[right] = i;
    if (i == 0) parent_internal[0] = -1;
}

// ---- refit --------------------------------------------------------------------------------------
__device__ __forceinline__ void load_child_box(int32_t ref, const float* aabbs, const uint32_t* sorted_prims, const float* node_boxes, float* lo, float* hi) {
    if (ref < 0) {
        const float* b = aabbs + 6 * (size_t)sorted_prims[~ref];
        for (int k = 0; k < 3; k++) { lo[k] = b[k]; hi[k] = b[k + 3]; }
    } else {
        // written by another thread during this launch: bypass L1 (a neighbouring box may have
        // pulled a stale copy of the line into this SM earlier)
        const float* b = node_boxes + 6 * (size_t)ref;
        for (int k = 0; k < 3; k++) { lo[k] = __ldcg(b + k); hi[k] = __ldcg(b + k + 3); }
    }
}

__global__ void k_refit(const float* __restrict__ aabbs, const uint32_t* __restrict__ sorted_prims, int n, const int32_t* __restrict__ children,
                        const int32_t* __restrict__ parent_internal, const int32_t* __restrict__ parent_leaf, uint32_t* flags, float* node_boxes,
                        float4* nodes, const uint8_t* __restrict__ prim_tags) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= n) return;
    int node = parent_leaf[leaf];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&flags[node], 1u) == 0u) return; // first arrival stops; the second sees both children done
        __threadfence();
        const int32_t r0 = children[2 * node], r1 = children[2 * node + 1];
        float lo0[3], hi0[3], lo1[3], hi1[3];
        load_child_box(r0, aabbs, sorted_prims, node_boxes, lo0, hi0);
        load_child_box(r1, aabbs, sorted_prims, node_boxes, lo1, hi1);
        float* nb = node_boxes + 6 * (size_t)node;
        for (int k = 0; k < 3; k++) {
            // volatile-ish publish: written before the fence + atomic of the next level
            nb[k] = fminf(lo0[k], lo1[k]);
            nb[k + 3] = fmaxf(hi0[k], hi1[k]);
        }
        // leaf refs in the final nodes name the ORIGINAL primitive index
        // (and, in bits 27..30, the caller's per-primitive tag: the traversal reads the primitive type from the ref)
        int32_t f0 = r0, f1 = r1;
        if (r0 < 0) { const uint32_t p = sorted_prims[~r0]; f0 = ~(int32_t)(p | (prim_tags ? (uint32_t)prim_tags[p] << 27 : 0u)); }
        if (r1 < 0) { const uint32_t p = sorted_prims[~r1]; f1 = ~(int32_t)(p | (prim_tags ? (uint32_t)prim_tags[p] << 27 : 0u)); }
        float4* out = nodes + 4 * (size_t)node;
        out[0] = make_float4(lo0[0], lo0[1], lo0[2], hi0[0]);
        out[1] = make_float4(hi0[1], hi0[2], lo1[0], lo1[1]);
        out[2] = make_float4(lo1[2], hi1[0], hi1[1], hi1[2]);
        out[3] = make_float4(__int_as_float(f0), __int_as_float(f1), 0.0f, 0.0f);
        node = parent_internal[node];
    }
}


// ---- 4-wide collapse ------------------------------------------------------------------------------
// The walk moves one node per dependent fetch, so a binary tree costs a fetch per level; with two levels per node the
// same boxes are tested in half the fetches (the tests of a node's four children are independent: ILP instead of a
// dependency).  Wide node i is built from binary node i: each child that is a leaf keeps its slot, each internal child
// is replaced by ITS two children.  Every binary node gets a wide node (one thread each, no ordering between them);
// only those on even levels below the root are ever reached.  128 bytes, SoA so a thread tests four boxes per float4:
//   [0] lo.x[4]  [1] lo.y[4]  [2] lo.z[4]  [3] hi.x[4]  [4] hi.y[4]  [5] hi.z[4]  [6] refs[4] (int bits)  [7] unused
// ref >= 0: wide node index; < 0: leaf (~(primitive | tag << 27)); kWideInvalid: empty slot.  The boxes are the binary
// tree's, bit for bit, so the walk's slab test sees the same numbers.
__global__ void k_collapse4(const float4* __restrict__ nodes, int n_internal, float4* __restrict__ wide) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_internal) return;
    float lo[4][3], hi[4][3];
    int32_t ref[4];
    int n = 0;
    auto take = [&](const float4* nd, int which) { // child `which` of binary node nd
        const float4 a = nd[0], b = nd[1], c = nd[2], e = nd[3];
        if (which == 0) { lo[n][0] = a.x; lo[n][1] = a.y; lo[n][2] = a.z; hi[n][0] = a.w; hi[n][1] = b.x; hi[n][2] = b.y; ref[n] = __float_as_int(e.x); }
        else { lo[n][0] = b.z; lo[n][1] = b.w; lo[n][2] = c.x; hi[n][0] = c.y; hi[n][1] = c.z; hi[n][2] = c.w; ref[n] = __float_as_int(e.y); }
        n++;
    };
    const float4* me = nodes + 4 * (size_t)i;
    const float4 e = me[3];
    const int32_t r[2] = {__float_as_int(e.x), __float_as_int(e.y)};
    for (int c = 0; c < 2; c++) {
        if (r[c] < 0) take(me, c);
        else { const float4* ch = nodes + 4 * (size_t)r[c]; take(ch, 0); take(ch, 1); }
    }
    for (; n < 4; n++) {
        for (int k = 0; k < 3; k++) { lo[n][k] = __builtin_inff(); hi[n][k] = -__builtin_inff(); }
        ref[n] = kWideInvalid;
    }
    float4* out = wide + 8 * (size_t)i;
    for (int k = 0; k < 3; k++) {
        out[k] = make_float4(lo[0][k], lo[1][k], lo[2][k], lo[3][k]);
        out[3 + k] = make_float4(hi[0][k], hi[1][k], hi[2][k], hi[3][k]);
    }
    out[6] = make_float4(__int_as_float(ref[0]), __int_as_float(ref[1]), __int_as_float(ref[2]), __int_as_float(ref[3]));
    out[7] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

// ---- the top of the tree, packed for shared memory --------------------------------------------------
// Breadth first from the root, `capacity` nodes at most: top[j] is a copy of a wide node whose child refs point INTO the
// array (kWideTopFlag | index) wherever the child made it in, and to the global array otherwise.  The walk kernels copy
// the array into shared memory once per block; every ray starts there.  One block; the frontier of a level is processed
// between two barriers.  *n_top = nodes written.
__global__ void k_build_top(const float4* __restrict__ wide, int32_t root, uint32_t capacity, float4* __restrict__ top, uint32_t* n_top) {
    extern __shared__ int32_t s_ids[]; // [capacity]: wide node index of top[j]
    __shared__ uint32_t s_count;
    if (threadIdx.x == 0) { s_ids[0] = root; s_count = 1u; }
    __syncthreads();
    uint32_t begin = 0;
    for (;;) {
        const uint32_t end = s_count < capacity ? s_count : capacity;
        __syncthreads();
        if (begin >= end) break;
        for (uint32_t j = begin + threadIdx.x; j < end; j += blockDim.x) {
            const float4* src = wide + 8 * (size_t)s_ids[j];
            float4 refs = src[6];
            int32_t r[4] = {__float_as_int(refs.x), __float_as_int(refs.y), __float_as_int(refs.z), __float_as_int(refs.w)};
            for (int k = 0; k < 4; k++) {
                if (r[k] < 0 || r[k] == kWideInvalid) continue;
                const uint32_t at = atomicAdd(&s_count, 1u);
                if (at < capacity) { s_ids[at] = r[k]; r[k] = (int32_t)(kWideTopFlag | at); }
            }
            float4* dst = top + 8 * (size_t)j;
            for (int k = 0; k < 6; k++) dst[k] = src[k];
            dst[6] = make_float4(__int_as_float(r[0]), __int_as_float(r[1]), __int_as_float(r[2]), __int_as_float(r[3]));
            dst[7] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
        begin = end;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_top = s_count < capacity ? s_count : capacity;
}

} // namespace

int lbvh_collapse(const float4* d_nodes, uint32_t n, float4* d_wide, float4* d_top, uint32_t top_capacity, uint32_t* d_n_top, void* stream_) {
    cudaStream_t s = (cudaStream_t)stream_;
    if (n < 2) return 0;
    const int n_internal = (int)n - 1;
    k_collapse4<<<(n_internal + kBlock - 1) / kBlock, kBlock, 0, s>>>(d_nodes, n_internal, d_wide);
    k_build_top<<<1, 256, top_capacity * sizeof(int32_t), s>>>(d_wide, 0, top_capacity, d_top, d_n_top);
    return 2;
}

size_t lbvh_scratch_bytes(uint32_t n) {
    const size_t n_warps = (n + kKeysPerWarp - 1) / kKeysPerWarp;
    size_t b = 0;
    b += 8 * sizeof(uint32_t);                 // bounds
    b += 4 * (size_t)n * sizeof(uint32_t);     // keys/vals ping-pong
    b += 256 * n_warps * sizeof(uint32_t);     // radix histogram
    b += 2 * (size_t)n * sizeof(int32_t);      // children
    b += 2 * (size_t)n * sizeof(int32_t);      // parents
    b += (size_t)n * sizeof(uint32_t);         // flags
    b += 6 * (size_t)n * sizeof(float);        // node boxes
    return b + 16 * 256; // every sub-allocation is rounded up to 256 bytes
}

int lbvh_build(const float* d_aabbs, uint32_t n, void* d_scratch, float4* d_nodes, uint32_t* d_sorted_prims, int32_t* root_ref, void* stream_,
               const uint8_t* d_prim_tags) {
    cudaStream_t s = (cudaStream_t)stream_;
    int launches = 0;
    if (n == 0) { *root_ref = INT32_MIN; return 0; }
    if (n == 1) { *root_ref = ~0; cudaMemsetAsync(d_sorted_prims, 0, sizeof(uint32_t), s); return 0; }
    const uint32_t n_warps = (n + kKeysPerWarp - 1) / kKeysPerWarp;
    char* p = (char*)d_scratch;
    auto take = [&](size_t bytes) { char* q = p; p += (bytes + 255) & ~(size_t)255; return (void*)q; };
    uint32_t* bounds = (uint32_t*)take(8 * sizeof(uint32_t));
    uint32_t* keys_a = (uint32_t*)take((size_t)n * 4);
    uint32_t* keys_b = (uint32_t*)take((size_t)n * 4);
    uint32_t* vals_a = (uint32_t*)take((size_t)n * 4);
    uint32_t* vals_b = (uint32_t*)take((size_t)n * 4);
    uint32_t* hist = (uint32_t*)take(256 * (size_t)n_warps * 4);
    int32_t* children = (int32_t*)take(2 * (size_t)n * 4);
    int32_t* parent_internal = (int32_t*)take((size_t)n * 4);
    int32_t* parent_leaf = (int32_t*)take((size_t)n * 4);
    uint32_t* flags = (uint32_t*)take((size_t)n * 4);
    float* node_boxes = (float*)take(6 * (size_t)n * 4);

    const int grid_n = (int)((n + kBlock - 1) / kBlock);
    // bounds: min slots start at all-ones, max slots at zero (ordered encoding)
    cudaMemsetAsync(bounds, 0xFF, 3 * sizeof(uint32_t), s);
    cudaMemsetAsync(bounds + 3, 0x00, 3 * sizeof(uint32_t), s);
    k_bounds<<<grid_n < 1184 ? grid_n : 1184, kBlock, 0, s>>>(d_aabbs, n, bounds); launches++;
    k_morton<<<grid_n, kBlock, 0, s>>>(d_aabbs, n, bounds, keys_a, vals_a); launches++;
    const int grid_w = (int)((n_warps + kWarpsPerBlock - 1) / kWarpsPerBlock);
    uint32_t *ki = keys_a, *ko = keys_b, *vi = vals_a, *vo = vals_b;
    for (int pass = 0; pass < 4; pass++) {
        const int shift = 8 * pass;
        k_radix_hist<<<grid_w, kBlock, 0, s>>>(ki, n, shift, hist, n_warps); launches++;
        k_scan_exclusive<<<1, 1024, 0, s>>>(hist, 256u * n_warps); launches++;
        k_radix_scatter<<<grid_w, kBlock, 0, s>>>(ki, vi, ko, vo, n, shift, hist, n_warps); launches++;
        uint32_t* t = ki; ki = ko; ko = t;
        t = vi; vi = vo; vo = t;
    }
    // after 4 passes the sorted data is back in keys_a / vals_a
    cudaMemcpyAsync(d_sorted_prims, vi, (size_t)n * 4, cudaMemcpyDeviceToDevice, s);
    k_karras<<<grid_n, kBlock, 0, s>>>(ki, (int)n, children, parent_internal, parent_leaf); launches++;
    cudaMemsetAsync(flags, 0, (size_t)n * 4, s);
    k_refit<<<grid_n, kBlock, 0, s>>>(d_aabbs, d_sorted_prims, (int)n, children, parent_internal, parent_leaf, flags, node_boxes, d_nodes, d_prim_tags); launches++;
    *root_ref = 0;
    return launches;
}

int lbvh_refit(const float* d_aabbs, uint32_t n, void* d_scratch, float4* d_nodes, const uint32_t* d_sorted_prims, void* stream_,
               const uint8_t* d_prim_tags) {
    // scratch layout must match lbvh_build (children / parents are reused)
    cudaStream_t s = (cudaStream_t)stream_;
    if (n < 2) return 0;
    const uint32_t n_warps = (n + kKeysPerWarp - 1) / kKeysPerWarp;
    char* p = (char*)d_scratch;
    auto take = [&](size_t bytes) { char* q = p; p += (bytes + 255) & ~(size_t)255; return (void*)q; };
    take(8 * sizeof(uint32_t));
    for (int i = 0; i < 4; i++) take((size_t)n * 4);
    take(256 * (size_t)n_warps * 4);
    int32_t* children = (int32_t*)take(2 * (size_t)n * 4);
    int32_t* parent_internal = (int32_t*)take((size_t)n * 4);
    int32_t* parent_leaf = (int32_t*)take((size_t)n * 4);
    uint32_t* flags = (uint32_t*)take((size_t)n * 4);
    float* node_boxes = (float*)take(6 * (size_t)n * 4);
    cudaMemsetAsync(flags, 0, (size_t)n * 4, s);
    k_refit<<<(int)((n + kBlock - 1) / kBlock), kBlock, 0, s>>>(d_aabbs, d_sorted_prims, (int)n, children, parent_internal, parent_leaf, flags, node_boxes, d_nodes, d_prim_tags);
    return 1;
}

} // namespace gprt
