// Titan/vec.h -- titan::Vec, the 3-vector of the public API (drop-in for the reference's include/Titan/vec.h:33-166).
//
// Host-only and header-only here: the B200 engine keeps structure-of-arrays doubles on the device and never
// ships Vec objects across the C ABI.  The arithmetic a user can observe is kept operation for operation:
//   norm()  = sqrt(pow(x,2) + pow(y,2) + pow(z,2))          (reference vec.h:149-151)
//   dot(a,b)= (a*b).sum() = (a0*b0 + a1*b1) + a2*b2          (reference src/vec.cu:39-41)
//   v / s   = three true divisions                           (reference vec.h:133-135)
// so builders (rest lengths, rotate) produce bit-identical doubles.
#ifndef TITAN_B200_VEC_H
#define TITAN_B200_VEC_H

#include <cmath>
#include <cstdio>
#include <ostream>
#include <vector>

namespace titan {

class Vec {
public:
    Vec() : data{0.0, 0.0, 0.0} {}
    Vec(double x, double y, double z) : data{x, y, z} {}
    Vec(const std::vector<double> &v) : data{v[0], v[1], v[2]} {}
    Vec(const Vec &) = default;
    Vec &operator=(const Vec &) = default;

    Vec &operator+=(const Vec &v) { for (int c = 0; c < 3; c++) data[c] += v.data[c]; return *this; }
    Vec &operator-=(const Vec &v) { for (int c = 0; c < 3; c++) data[c] -= v.data[c]; return *this; }
    Vec operator-() const { return Vec(-data[0], -data[1], -data[2]); }

    // out-of-range access reports and aliases component 0, like the reference (vec.h:83-99)
    double &operator[](int n) { return data[check(n)]; }
    const double &operator[](int n) const { return data[check(n)]; }

    friend Vec operator+(const Vec &a, const Vec &b) { return Vec(a.data[0] + b.data[0], a.data[1] + b.data[1], a.data[2] + b.data[2]); }
    friend Vec operator-(const Vec &a, const Vec &b) { return Vec(a.data[0] - b.data[0], a.data[1] - b.data[1], a.data[2] - b.data[2]); }
    friend Vec operator*(double s, const Vec &v) { return Vec(v.data[0] * s, v.data[1] * s, v.data[2] * s); }
    friend Vec operator*(const Vec &v, double s) { return s * v; }
    friend Vec operator*(const Vec &a, const Vec &b) { return Vec(a.data[0] * b.data[0], a.data[1] * b.data[1], a.data[2] * b.data[2]); }
    friend Vec operator/(const Vec &v, double s) { return Vec(v.data[0] / s, v.data[1] / s, v.data[2] / s); }
    friend Vec operator/(const Vec &a, const Vec &b) { return Vec(a.data[0] / b.data[0], a.data[1] / b.data[1], a.data[2] / b.data[2]); }
    friend bool operator==(const Vec &a, const Vec &b) { return a.data[0] == b.data[0] && a.data[1] == b.data[1] && a.data[2] == b.data[2]; }
    friend std::ostream &operator<<(std::ostream &os, const Vec &v) { return os << "(" << v.data[0] << ", " << v.data[1] << ", " << v.data[2] << ")"; }

    void print() const { std::printf("(%3f, %3f, %3f)\n", data[0], data[1], data[2]); }
    double norm() const { return std::sqrt(std::pow(data[0], 2) + std::pow(data[1], 2) + std::pow(data[2], 2)); }
    double sum() const { return data[0] + data[1] + data[2]; }
    Vec normalize() const { return *this / norm(); }

private:
    static int check(int n) {
        if (n < 0 || n >= 3) { std::printf("%s\n", "Out of bounds!"); return 0; }
        return n;
    }
    double data[3];
};

inline double dot(const Vec &a, const Vec &b) { return (a * b).sum(); }
inline Vec cross(const Vec &a, const Vec &b) {
    return Vec(a[1] * b[2] - a[2] * b[1], b[0] * a[2] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]);
}

}  // namespace titan

#endif
