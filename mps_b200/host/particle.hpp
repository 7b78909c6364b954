// particle.hpp — host mirror of the reference's particle record (reference src/particle.hpp:5-46).
// Same names, field meaning and enum values; this is what `MPS::particles` holds and what
// `Exporter` reads. On the B200 path the authoritative state lives in HBM as SoA; this
// AoS record is filled from it by MPS::downloadParticles() on export steps only.
#pragma once

#include "vec3.hpp"

#include <vector>

enum class ParticleType { Ghost = -1, Fluid = 0, Wall = 1, DummyWall = 2 };

class Neighbor {
public:
    int id;
    double distance;
    Neighbor(int id, double distance) : id(id), distance(distance) {}
};

enum class BoundaryCondition { Ignored, Surface, Inner };

class Particle {
public:
    int id;
    ParticleType type;
    Eigen::Vector3d position, velocity;
    Eigen::Vector3d acceleration = Eigen::Vector3d::Zero();
    // The reference leaves pressure / boundaryCondition / isDirichletBoundaryConnected uninitialised
    // (SURVEY B-1); they are defined here: 0, Ignored, false.
    double pressure = 0, minimumPressure = 0;
    double density;
    double numberDensityRatio = 0;

    // Kept for source compatibility. NOT filled by the B200 path (the lists live in HBM as sliced ELL);
    // MPS::downloadNeighbors() materialises them on request.
    std::vector<Neighbor> neighbors;

    BoundaryCondition boundaryCondition = BoundaryCondition::Ignored;
    bool isDirichletBoundaryConnected  = false;

    Particle(
        const int& id,
        const ParticleType& type,
        const Eigen::Vector3d& position,
        const Eigen::Vector3d& velocity,
        const double& density
    )
        : id(id), type(type), position(position), velocity(velocity), density(density) {}

    // 1/density for Fluid, 0 for Wall / DummyWall / Ghost (reference src/particle.cpp:19-29;
    // the reference has no return for an invalid enum — here that is 0 as well).
    double inverseDensity() const { return type == ParticleType::Fluid ? 1.0 / density : 0.0; }
};
