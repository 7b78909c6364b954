// generator.hpp — the lattice of the reference's input generators (reference inputs/dambreak/main.cpp:26-80,
// inputs/dambreak_fine/main.cpp, inputs/hydrostatic/main.cpp: same program, different constants), as a function:
// an open-top tank [0, domainX] x [0, domainY] with `wallLayer` wall and `dummyWallLayer` dummy-wall layers on the
// bottom and both sides, and a fluid box. Particles sit at ((ix + 0.5) l0, (iy + 0.5) l0, 0), ids are dense in
// loop order (x outer, y inner), classification by the inclusive-with-tolerance box test of the reference
// (eps = 0.01 l0). Positions keep full double precision (the reference's CSV rounds them to 6 decimals).
#pragma once

#include "particle.hpp"

#include <cmath>
#include <stdexcept>
#include <vector>

namespace generator {

struct Tank {
    double l0;
    double domainX[2], domainY[2]; // inner (empty) region of the tank
    double fluidX[2], fluidY[2];
    int wallLayer      = 1;
    int dummyWallLayer = 2;
};

inline std::vector<Particle> lattice(const Tank& t, double density) {
    const double l0 = t.l0, eps = 0.01 * l0;
    auto integral = [&](double len) { return !(std::fabs(len / l0 - std::round(len / l0)) > eps); }; // the reference's test, eps = 0.01 l0 (sic)
    if (!integral(t.fluidX[1] - t.fluidX[0]) || !integral(t.fluidY[1] - t.fluidY[0]))
        throw std::runtime_error("fluid range is not divisible by the particle distance"); // the reference's checkFluidRange
    auto inside = [&](double x, double y, double x0, double x1, double y0, double y1) {
        return x0 - eps < x && x < x1 + eps && y0 - eps < y && y < y1 + eps;
    };
    const int all     = t.dummyWallLayer + t.wallLayer;
    const int ixBegin = (int) std::round(t.domainX[0] / l0) - all, ixEnd = (int) std::round(t.domainX[1] / l0) + all;
    const int iyBegin = (int) std::round(t.domainY[0] / l0) - all, iyEnd = (int) std::round(t.domainY[1] / l0) + all;
    const double xMax = t.domainX[1], yMax = t.domainY[1];
    std::vector<Particle> ps;
    for (int ix = ixBegin; ix <= ixEnd; ix++)
        for (int iy = iyBegin; iy <= iyEnd; iy++) {
            const double x = (ix + 0.5) * l0, y = (iy + 0.5) * l0;
            ParticleType type = ParticleType::Ghost;
            if (inside(x, y, -all * l0, xMax + all * l0, -all * l0, yMax)) type = ParticleType::DummyWall;
            if (inside(x, y, -t.wallLayer * l0, xMax + t.wallLayer * l0, -t.wallLayer * l0, yMax)) type = ParticleType::Wall;
            if (inside(x, y, 0.0, xMax, 0.0, yMax)) type = ParticleType::Ghost;
            if (inside(x, y, t.fluidX[0], t.fluidX[1], t.fluidY[0], t.fluidY[1])) type = ParticleType::Fluid;
            if (type != ParticleType::Ghost)
                ps.push_back(Particle((int) ps.size(), type, Eigen::Vector3d(x, y, 0.0), Eigen::Vector3d(0.0, 0.0, 0.0), density));
        }
    return ps;
}

} // namespace generator
