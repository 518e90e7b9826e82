/*
 * prt_math.h — FP64 vector arithmetic shared by the host flattener and the CUDA kernels.
 *
 * Every operation is spelled in the order the reference's vendored glm evaluates it
 * (ThirdPartyLibraries/glm/glm/detail/func_geometric.inl:52-53 dot, :79-82 cross, :104 normalize,
 * func_exponential.inl:138 inversesqrt = 1/sqrt) so that host (g++ -ffp-contract=off) and device
 * (nvcc -fmad=false) produce the same bits as the reference's x86-64 build, which contains no FMA.
 */
#ifndef PRT_MATH_H
#define PRT_MATH_H

#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define PRT_HD __host__ __device__ __forceinline__
#else
#define PRT_HD inline
#endif

namespace prt {

struct d3 {
    double x, y, z;
};
struct d2 {
    double x, y;
};

PRT_HD d3 mk3(double x, double y, double z) { d3 r; r.x = x; r.y = y; r.z = z; return r; }
PRT_HD d2 mk2(double x, double y) { d2 r; r.x = x; r.y = y; return r; }
PRT_HD d3 operator+(d3 a, d3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
PRT_HD d3 operator-(d3 a, d3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
PRT_HD d3 operator-(d3 a) { return mk3(-a.x, -a.y, -a.z); }
PRT_HD d3 operator*(d3 a, double s) { return mk3(a.x * s, a.y * s, a.z * s); }
PRT_HD d3 operator*(double s, d3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
PRT_HD d3 operator*(d3 a, d3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
PRT_HD d3 operator/(d3 a, double s) { return mk3(a.x / s, a.y / s, a.z / s); }
PRT_HD d2 operator+(d2 a, d2 b) { return mk2(a.x + b.x, a.y + b.y); }
PRT_HD d2 operator-(d2 a, d2 b) { return mk2(a.x - b.x, a.y - b.y); }
PRT_HD d2 operator*(double s, d2 a) { return mk2(s * a.x, s * a.y); }
PRT_HD double comp(d3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

/* glm::dot(dvec3): tmp = a*b; (tmp.x + tmp.y) + tmp.z */
PRT_HD double dot(d3 a, d3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
/* glm::cross */
PRT_HD d3 cross(d3 x, d3 y) {
    return mk3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
PRT_HD double length(d3 a) { return sqrt(dot(a, a)); }
/* glm::normalize: v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x) */
PRT_HD d3 normalize(d3 a) { return a * (1.0 / sqrt(dot(a, a))); }
PRT_HD bool anynan(d3 a) { return a.x != a.x || a.y != a.y || a.z != a.z; }

/* Source/RandomNumberGenerator.h:10-14 */
#define PRT_PI 3.14159265358979323846
#define PRT_INV_PI 0.31830988618379067154
#define PRT_INV_2PI 0.15915494309189533577
#define PRT_PI_OVER_2 1.57079632679489661923
#define PRT_PI_OVER_4 0.78539816339744830961

}  // namespace prt
#endif
