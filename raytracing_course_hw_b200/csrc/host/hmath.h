// Host-side float math for the loaders and the BVH builder.  Operation order follows the
// reference's Vec3 (hw8/src/include/vec3.h:32-80) so that loaded geometry is bit-identical:
// note normalize() = float(1.0 / double(len)) * v and cross() = b x a (negated).
#pragma once
#include <cmath>
#include <cstring>

namespace b200rt_host {

struct V3 {
    float x = 0, y = 0, z = 0;
    V3() {}
    V3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    V3 operator+(const V3 &o) const { return {x + o.x, y + o.y, z + o.z}; }
    V3 operator-(const V3 &o) const { return {x - o.x, y - o.y, z - o.z}; }
    float dot(const V3 &o) const { return x * o.x + y * o.y + z * o.z; }
    V3 cross_ref(const V3 &o) const { return {z * o.y - y * o.z, x * o.z - z * o.x, y * o.x - x * o.y}; }
    float len() const { return std::sqrt(dot(*this)); }
    V3 normalized() const {
        float k = (float)(1. / len());
        return {k * x, k * y, k * z};
    }
};

// Row-major 4x4 affine transform; mirrors the arithmetic of the reference's `Transition`
// (hw8/src/include/transition.h:7-139): compose() accumulates k = 0..3 into a zeroed cell,
// apply() is m[i][0] x + m[i][1] y + m[i][2] z + m[i][3].
struct Mat4 {
    float m[4][4];
    Mat4() {
        memset(m, 0, sizeof(m));
        m[0][0] = m[1][1] = m[2][2] = m[3][3] = 1;
    }
    static Mat4 scale(const V3 &s) {
        Mat4 r;
        r.m[0][0] = s.x; r.m[1][1] = s.y; r.m[2][2] = s.z;
        return r;
    }
    static Mat4 translation(const V3 &t) {
        Mat4 r;
        r.m[0][3] = t.x; r.m[1][3] = t.y; r.m[2][3] = t.z;
        return r;
    }
    // Rotation from a glTF quaternion (x, y, z, w); element formulas of transition.h:35-43.
    static Mat4 rotation(float qx, float qy, float qz, float qw) {
        Mat4 r;
        r.m[0][0] = 2 * (qw * qw + qx * qx) - 1; r.m[0][1] = 2 * (qx * qy - qw * qz);     r.m[0][2] = 2 * (qx * qz + qw * qy);
        r.m[1][0] = 2 * (qx * qy + qw * qz);     r.m[1][1] = 2 * (qw * qw + qy * qy) - 1; r.m[1][2] = 2 * (qy * qz - qw * qx);
        r.m[2][0] = 2 * (qx * qz - qw * qy);     r.m[2][1] = 2 * (qy * qz + qw * qx);     r.m[2][2] = 2 * (qw * qw + qz * qz) - 1;
        return r;
    }
    Mat4 compose(const Mat4 &o) const {
        Mat4 r;
        memset(r.m, 0, sizeof(r.m));
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++)
                for (int k = 0; k < 4; k++) r.m[i][j] += m[i][k] * o.m[k][j];
        return r;
    }
    V3 apply(const V3 &p) const {
        float r[3];
        for (int i = 0; i < 3; i++) r[i] = m[i][0] * p.x + m[i][1] * p.y + m[i][2] * p.z + m[i][3];
        return {r[0], r[1], r[2]};
    }
    Mat4 transposed() const {
        Mat4 r;
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) r.m[j][i] = m[i][j];
        return r;
    }
    // General inverse by the adjugate: inv[j][i] = (-1)^(i+j) * minor(i, j) / det.
    // (The reference expands the same cofactors term by term, transition.h:77-128; summation
    // order inside a 3x3 minor may differ in the last ulp, which only touches normals that are
    // re-normalised afterwards.)
    Mat4 inverted() const {
        auto minor3 = [&](int ri, int cj) {
            int r[3], c[3], a = 0, b = 0;
            for (int i = 0; i < 4; i++) if (i != ri) r[a++] = i;
            for (int j = 0; j < 4; j++) if (j != cj) c[b++] = j;
            return m[r[0]][c[0]] * (m[r[1]][c[1]] * m[r[2]][c[2]] - m[r[1]][c[2]] * m[r[2]][c[1]])
                 - m[r[0]][c[1]] * (m[r[1]][c[0]] * m[r[2]][c[2]] - m[r[1]][c[2]] * m[r[2]][c[0]])
                 + m[r[0]][c[2]] * (m[r[1]][c[0]] * m[r[2]][c[1]] - m[r[1]][c[1]] * m[r[2]][c[0]]);
        };
        float cof[4][4];
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) cof[i][j] = (((i + j) & 1) ? -1.f : 1.f) * minor3(i, j);
        float det = m[0][0] * cof[0][0] + m[0][1] * cof[0][1] + m[0][2] * cof[0][2] + m[0][3] * cof[0][3];
        float inv_det = (float)(1.0 / det);
        Mat4 r;
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) r.m[j][i] = cof[i][j] * inv_det;
        return r;
    }
};

}  // namespace b200rt_host
