// mathx.h -- the few DirectXMath functions the reference's Renderer calls host-side
// (src/Renderer.cpp:255-265, 289-325, 742-782), restated.  DirectXMath is not vendored in the
// reference.  Row-vector convention: v' = v * M, translation in row 3, row-major storage.
// Everything is evaluated in double and rounded to float once when stored into the ABI structs.
// The matrix functions are __host__ __device__ so that the per-frame transform update can run on the
// GPU (csrc/transforms.cu) with exactly the host's arithmetic (double, no FMA contraction either side).
#pragma once
#include <cmath>
#include <cstring>

#if defined(__CUDACC__)
#define GPRT_MX __host__ __device__ inline
#else
#define GPRT_MX inline
#endif

namespace gprt_host {

struct Mat4 {
    double m[4][4];
};

GPRT_MX Mat4 mat_identity() {
    Mat4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) r.m[i][j] = (i == j) ? 1.0 : 0.0;
    return r;
}
GPRT_MX Mat4 mat_mul(const Mat4& a, const Mat4& b) { // XMMatrixMultiply(a, b) == a * b
    Mat4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            double s = 0;
            for (int k = 0; k < 4; k++) s += a.m[i][k] * b.m[k][j];
            r.m[i][j] = s;
        }
    return r;
}
GPRT_MX Mat4 mat_scaling(double x, double y, double z) { // XMMatrixScaling
    Mat4 r = mat_identity();
    r.m[0][0] = x; r.m[1][1] = y; r.m[2][2] = z;
    return r;
}
GPRT_MX Mat4 mat_translation(double x, double y, double z) { // XMMatrixTranslationFromVector
    Mat4 r = mat_identity();
    r.m[3][0] = x; r.m[3][1] = y; r.m[3][2] = z;
    return r;
}
inline Mat4 mat_rotation_y(double a) { // XMMatrixRotationY
    Mat4 r = mat_identity();
    const double c = std::cos(a), s = std::sin(a);
    r.m[0][0] = c; r.m[0][2] = -s;
    r.m[2][0] = s; r.m[2][2] = c;
    return r;
}
inline void vec3_transform(const double v[3], const Mat4& m, double out[4]) { // XMVector3Transform: (v,1) * M
    double r[4];
    for (int j = 0; j < 4; j++) r[j] = v[0] * m.m[0][j] + v[1] * m.m[1][j] + v[2] * m.m[2][j] + m.m[3][j];
    for (int j = 0; j < 4; j++) out[j] = r[j];
}
inline void vec3_normalize(double v[3]) {
    const double l = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    v[0] /= l; v[1] /= l; v[2] /= l;
}
inline void vec3_cross(const double a[3], const double b[3], double o[3]) {
    const double r0 = a[1] * b[2] - a[2] * b[1], r1 = a[2] * b[0] - a[0] * b[2], r2 = a[0] * b[1] - a[1] * b[0];
    o[0] = r0; o[1] = r1; o[2] = r2;
}
inline double vec3_dot(const double a[3], const double b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

inline Mat4 mat_look_at_lh(const double eye[3], const double at[3], const double up[3]) { // XMMatrixLookAtLH
    double z[3] = {at[0] - eye[0], at[1] - eye[1], at[2] - eye[2]};
    vec3_normalize(z);
    double x[3];
    vec3_cross(up, z, x);
    vec3_normalize(x);
    double y[3];
    vec3_cross(z, x, y);
    Mat4 r = mat_identity();
    for (int i = 0; i < 3; i++) { r.m[i][0] = x[i]; r.m[i][1] = y[i]; r.m[i][2] = z[i]; }
    r.m[3][0] = -vec3_dot(x, eye); r.m[3][1] = -vec3_dot(y, eye); r.m[3][2] = -vec3_dot(z, eye);
    return r;
}
inline Mat4 mat_perspective_fov_lh(double fov_y, double aspect, double zn, double zf) { // XMMatrixPerspectiveFovLH
    Mat4 r;
    std::memset(&r, 0, sizeof r);
    const double h = 1.0 / std::tan(fov_y * 0.5), w = h / aspect, q = zf / (zf - zn);
    r.m[0][0] = w; r.m[1][1] = h; r.m[2][2] = q; r.m[2][3] = 1.0; r.m[3][2] = -q * zn;
    return r;
}
// Every row and column index below is a compile-time constant once the loops are unrolled (the pivot search keeps the
// pivot's magnitude in a scalar and the row exchange tests `p == r` for each candidate row), so that the device build keeps
// the 4 x 8 work array in registers and drops the columns nothing reads again; a run-time row index put it in local memory
// (0.24 ms per frame for a million primitives, a fixed cost per GPU that one eighth of an 8K frame could not amortise).
// The operations and their order are those of the textbook loop: same bits on the host and on the device.
#if defined(__CUDACC__)
#define GPRT_MX_UNROLL _Pragma("unroll")
#else
#define GPRT_MX_UNROLL
#endif
GPRT_MX Mat4 mat_inverse(const Mat4& a) { // XMMatrixInverse (general 4x4, Gauss-Jordan with partial pivoting)
    double w[4][8];
    GPRT_MX_UNROLL
    for (int i = 0; i < 4; i++) {
        GPRT_MX_UNROLL
        for (int j = 0; j < 4; j++) { w[i][j] = a.m[i][j]; w[i][j + 4] = (i == j) ? 1.0 : 0.0; }
    }
    GPRT_MX_UNROLL
    for (int c = 0; c < 4; c++) {
        int p = c;
        double best = fabs(w[c][c]);
        GPRT_MX_UNROLL
        for (int r = c + 1; r < 4; r++) {
            const double v = fabs(w[r][c]);
            if (v > best) { best = v; p = r; }
        }
        GPRT_MX_UNROLL
        for (int r = c + 1; r < 4; r++)
            if (p == r) {
                GPRT_MX_UNROLL
                for (int j = 0; j < 8; j++) { const double t = w[c][j]; w[c][j] = w[r][j]; w[r][j] = t; }
            }
        const double d = w[c][c];
        GPRT_MX_UNROLL
        for (int j = 0; j < 8; j++) w[c][j] /= d;
        GPRT_MX_UNROLL
        for (int r = 0; r < 4; r++)
            if (r != c) {
                const double f = w[r][c];
                if (f != 0.0) {
                    GPRT_MX_UNROLL
                    for (int j = 0; j < 8; j++) w[r][j] -= f * w[c][j];
                }
            }
    }
    Mat4 r;
    GPRT_MX_UNROLL
    for (int i = 0; i < 4; i++) {
        GPRT_MX_UNROLL
        for (int j = 0; j < 4; j++) r.m[i][j] = w[i][j + 4];
    }
    return r;
}
GPRT_MX void mat_store(const Mat4& a, float out[16]) { // XMStoreFloat4x4 (row-major)
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) out[i * 4 + j] = (float)a.m[i][j];
}
inline void mat_store_3x4(const Mat4& a, float out[3][4]) { // XMStoreFloat3x4: transposed, column-vector form
    for (int i = 0; i < 3; i++) for (int j = 0; j < 4; j++) out[i][j] = (float)a.m[j][i];
}

} // namespace gprt_host
