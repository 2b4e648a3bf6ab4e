#pragma once
// Host-side float3 / 3x3 algebra of the public scene API (reference: src/core/crt_vector.h:7-137,
// crt_matrix.h:7-80, crt_transform.h:8-10).  The host only needs the operations that construct kernel inputs
// (face normals, smooth vertex normals, camera constants); ray math lives in the CUDA kernels, and there is no
// CPU render path.  Every operation is a single IEEE binary32 operation per step, evaluated left to right, so the
// values uploaded to the GPU are bit-identical to the ones the reference computes (build with -ffp-contract=off).

#include <cmath>

namespace crt {

struct Vector {
    union {
        struct {
            float x, y, z;
        };
        float data[3];
    };

    constexpr Vector() : x(0.0f), y(0.0f), z(0.0f) {}
    constexpr Vector(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}

    constexpr float dot(const Vector &o) const { return x * o.x + y * o.y + z * o.z; }
    constexpr float length_squared() const { return dot(*this); }
    float length() const { return std::sqrt(length_squared()); }

    constexpr Vector cross(const Vector &o) const { return { y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x }; }

    friend constexpr Vector operator+(const Vector &a, const Vector &b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
    friend constexpr Vector operator-(const Vector &a, const Vector &b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
    friend constexpr Vector operator-(const Vector &a) { return { -a.x, -a.y, -a.z }; }
    friend constexpr Vector operator*(const Vector &a, float s) { return { a.x * s, a.y * s, a.z * s }; }
    friend constexpr Vector operator/(const Vector &a, float s) { return { a.x / s, a.y / s, a.z / s }; }
    constexpr Vector &operator+=(const Vector &o) { return *this = *this + o; }
    constexpr Vector &operator-=(const Vector &o) { return *this = *this - o; }
    constexpr Vector &operator*=(float s) { return *this = *this * s; }
    constexpr Vector &operator/=(float s) { return *this = *this / s; }

    Vector normalized() const { return *this / length(); }
    Vector &normalize() { return *this = normalized(); }
};

// Row-major 3x3; vectors multiply from the left as ROW vectors (crt_matrix.h:66-74).
struct Matrix {
    float data[3][3]{};

    static constexpr Matrix identity() {
        Matrix m{};
        m.data[0][0] = m.data[1][1] = m.data[2][2] = 1.0f;
        return m;
    }
    static constexpr Matrix from_axes(const Vector &right, const Vector &up, const Vector &forward) {
        Matrix m{};
        for (int c = 0; c < 3; ++c) {
            m.data[0][c] = right.data[c];
            m.data[1][c] = up.data[c];
            m.data[2][c] = forward.data[c];
        }
        return m;
    }
    static Matrix rotation_x(float a) {
        Matrix m = identity();
        m.data[1][1] = std::cos(a), m.data[1][2] = std::sin(a);
        m.data[2][1] = -std::sin(a), m.data[2][2] = std::cos(a);
        return m;
    }
    static Matrix rotation_y(float a) {
        Matrix m = identity();
        m.data[0][0] = std::cos(a), m.data[0][2] = -std::sin(a);
        m.data[2][0] = std::sin(a), m.data[2][2] = std::cos(a);
        return m;
    }
    static Matrix rotation_z(float a) {
        Matrix m = identity();
        m.data[0][0] = std::cos(a), m.data[0][1] = std::sin(a);
        m.data[1][0] = -std::sin(a), m.data[1][1] = std::cos(a);
        return m;
    }
    // A true matrix product.  NOTE: the reference's Matrix::operator*= (crt_matrix.h:45-54) accumulates into
    // the operand it is still reading and is not a matrix product; nothing on the render path calls it
    // (SURVEY.md Appendix A item 20), so it is deliberately not reproduced.
    friend constexpr Matrix operator*(const Matrix &a, const Matrix &b) {
        Matrix r{};
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                for (int k = 0; k < 3; ++k)
                    r.data[i][j] += a.data[i][k] * b.data[k][j];
        return r;
    }
    friend constexpr Vector operator*(const Vector &v, const Matrix &m) {
        Vector r{};
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                r.data[i] += v.data[j] * m.data[j][i];
        return r;
    }
};

struct Transform {
    Vector location{};
    Matrix rotation{ Matrix::identity() };

    constexpr void translate_world(const Vector &t) { location += t; }
    constexpr void translate_local(const Vector &t) { location += t * rotation; }
    void rotate_x(float a) { rotation = rotation * Matrix::rotation_x(a); }
    void rotate_y(float a) { rotation = rotation * Matrix::rotation_y(a); }
    void rotate_z(float a) { rotation = rotation * Matrix::rotation_z(a); }
    void rotate_y_around(float a, const Vector &anchor) {
        const Matrix r = Matrix::rotation_y(a);
        rotation = rotation * r;
        location = (location - anchor) * r + anchor;
    }
    void rotate_x_around(float a, const Vector &anchor) {
        const Matrix r = Matrix::rotation_x(a);
        rotation = rotation * r;
        location = (location - anchor) * r + anchor;
    }
};

} // namespace crt
