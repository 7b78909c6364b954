// vec3.hpp — the three-component vector type of the host mirror.
//
// The reference's public records hold Eigen::Vector3d (reference src/particle.hpp:24-25,
// src/settings.hpp:52). The host mirror needs none of Eigen's arithmetic (all arithmetic of the path
// runs on the GPU), only a value type with the accessors callers of `Particle` / `Settings` use:
// x() y() z(), operator[], data(). With -DMPSB_HOST_USE_EIGEN the real Eigen type is used instead
// (a maintainer dropping these files into the reference tree).
#pragma once

#ifdef MPSB_HOST_USE_EIGEN
#include <Eigen/Dense>
#else
#include <cmath>
#include <cstddef>

namespace Eigen {
class Vector3d {
public:
    Vector3d() : c_{0.0, 0.0, 0.0} {}
    Vector3d(double x, double y, double z) : c_{x, y, z} {}
    static Vector3d Zero() { return Vector3d(); }
    double& x() { return c_[0]; }
    double& y() { return c_[1]; }
    double& z() { return c_[2]; }
    const double& x() const { return c_[0]; }
    const double& y() const { return c_[1]; }
    const double& z() const { return c_[2]; }
    double& operator[](std::size_t i) { return c_[i]; }
    const double& operator[](std::size_t i) const { return c_[i]; }
    double* data() { return c_; }
    const double* data() const { return c_; }
    // Eigen's reduction order for 3-vectors: (c0 + c1) + c2
    double squaredNorm() const { return (c_[0] * c_[0] + c_[1] * c_[1]) + c_[2] * c_[2]; }
    double norm() const { return std::sqrt(squaredNorm()); }

private:
    double c_[3];
};
} // namespace Eigen
#endif
