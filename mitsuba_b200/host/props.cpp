// props.cpp -- Transform factories (reference: src/libcore/transform.cpp:25-214).
#include "props.h"

#include <cmath>

namespace mtsbhost {

static Vec3 vnormalize(const Vec3 &v) {
    float l = std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z);
    if (l == 0) throw std::runtime_error("normalize(): zero-length vector");
    return Vec3{v.x / l, v.y / l, v.z / l};
}
static Vec3 vcross(const Vec3 &a, const Vec3 &b) { return Vec3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

Transform Transform::translate(const Vec3 &v) {
    Transform t;
    t.m[3] = v.x; t.m[7] = v.y; t.m[11] = v.z;
    return t;
}

Transform Transform::scale(const Vec3 &v) {
    Transform t;
    t.m[0] = v.x; t.m[5] = v.y; t.m[10] = v.z;
    return t;
}

// transform.cpp:65-97 (angle in degrees; Rodrigues form written per matrix entry as there)
Transform Transform::rotate(const Vec3 &axis, float angleDeg) {
    Vec3 a = vnormalize(axis);
    float rad = angleDeg * (float) (M_PI / 180.0);
    float s = std::sin(rad), c = std::cos(rad), k = 1.0f - c;
    Transform t;
    t.m[0] = a.x * a.x + (1.0f - a.x * a.x) * c;  t.m[1] = a.x * a.y * k - a.z * s;  t.m[2] = a.x * a.z * k + a.y * s;
    t.m[4] = a.x * a.y * k + a.z * s;  t.m[5] = a.y * a.y + (1.0f - a.y * a.y) * c;  t.m[6] = a.y * a.z * k - a.x * s;
    t.m[8] = a.x * a.z * k - a.y * s;  t.m[9] = a.y * a.z * k + a.x * s;  t.m[10] = a.z * a.z + (1.0f - a.z * a.z) * c;
    return t;
}

// transform.cpp:191-214: columns = (left, newUp, dir, p); left = up x dir
Transform Transform::lookAt(const Vec3 &p, const Vec3 &t, const Vec3 &up) {
    Vec3 dir = vnormalize(Vec3{t.x - p.x, t.y - p.y, t.z - p.z});
    Vec3 left = vnormalize(vcross(up, dir));
    Vec3 nup = vcross(dir, left);
    Transform r;
    r.m[0] = left.x; r.m[1] = nup.x; r.m[2] = dir.x;  r.m[3] = p.x;
    r.m[4] = left.y; r.m[5] = nup.y; r.m[6] = dir.y;  r.m[7] = p.y;
    r.m[8] = left.z; r.m[9] = nup.z; r.m[10] = dir.z; r.m[11] = p.z;
    return r;
}

// general 4x4 inverse by Gauss-Jordan with partial pivoting (matrix.h: Matrix::invert)
Transform Transform::inverse() const {
    double a[4][8];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) { a[i][j] = m[4 * i + j]; a[i][4 + j] = (i == j) ? 1.0 : 0.0; }
    for (int c = 0; c < 4; ++c) {
        int piv = c;
        for (int r = c + 1; r < 4; ++r) if (std::fabs(a[r][c]) > std::fabs(a[piv][c])) piv = r;
        if (a[piv][c] == 0.0) throw std::runtime_error("Transform::inverse(): singular matrix");
        if (piv != c) for (int j = 0; j < 8; ++j) std::swap(a[piv][j], a[c][j]);
        double inv = 1.0 / a[c][c];
        for (int j = 0; j < 8; ++j) a[c][j] *= inv;
        for (int r = 0; r < 4; ++r) {
            if (r == c) continue;
            double f = a[r][c];
            if (f != 0.0) for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j];
        }
    }
    Transform t;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) t.m[4 * i + j] = (float) a[i][4 + j];
    return t;
}

// Transform::operator()(Normal): n' = (M^-1)^T n
Vec3 Transform::normal(const Vec3 &n) const {
    Transform inv = inverse();
    return Vec3{inv.m[0] * n.x + inv.m[4] * n.y + inv.m[8] * n.z, inv.m[1] * n.x + inv.m[5] * n.y + inv.m[9] * n.z,
                inv.m[2] * n.x + inv.m[6] * n.y + inv.m[10] * n.z};
}

}  // namespace mtsbhost
