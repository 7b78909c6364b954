// state_io.hpp — binary particle state: the reference's CSV carries 5–6 decimals, far too
// coarse for bit-faithful restarts or for generated cases whose particle distance is below 1e-4. Layout
// (little-endian, no padding): char magic[8] = "MPSBST1\0"; int64 n; double time; then SoA arrays
// int32 type[n], int32 boundaryCondition[n], double position[3n], velocity[3n], pressure[n],
// numberDensityRatio[n]. Everything a step depends on (type, position, velocity — and, for the higher-order
// source term, the previous step's ratio and flags) is in it, so "run k steps, save, load, run m steps" equals
// "run k + m steps" bit for bit.
#pragma once

#include "particle.hpp"

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <filesystem>
#include <vector>

namespace state_io {

inline bool write(const std::filesystem::path& path, double time, const std::vector<Particle>& ps) {
    std::FILE* f = std::fopen(path.string().c_str(), "wb");
    if (!f) return false;
    const std::int64_t n = (std::int64_t) ps.size();
    const char magic[8]  = {'M', 'P', 'S', 'B', 'S', 'T', '1', '\0'};
    std::vector<std::int32_t> type((std::size_t) n), bc((std::size_t) n);
    std::vector<double> pos(3 * (std::size_t) n), vel(3 * (std::size_t) n), p((std::size_t) n), ratio((std::size_t) n);
    for (std::size_t i = 0; i < (std::size_t) n; i++) {
        type[i] = static_cast<std::int32_t>(ps[i].type);
        bc[i]   = static_cast<std::int32_t>(ps[i].boundaryCondition);
        for (std::size_t k = 0; k < 3; k++) {
            pos[3 * i + k] = ps[i].position[k];
            vel[3 * i + k] = ps[i].velocity[k];
        }
        p[i]     = ps[i].pressure;
        ratio[i] = ps[i].numberDensityRatio;
    }
    bool ok = std::fwrite(magic, 1, 8, f) == 8 && std::fwrite(&n, sizeof n, 1, f) == 1 && std::fwrite(&time, sizeof time, 1, f) == 1;
    auto put = [&](const void* d, std::size_t bytes) { ok = ok && (bytes == 0 || std::fwrite(d, 1, bytes, f) == bytes); };
    put(type.data(), 4 * type.size());
    put(bc.data(), 4 * bc.size());
    put(pos.data(), 8 * pos.size());
    put(vel.data(), 8 * vel.size());
    put(p.data(), 8 * p.size());
    put(ratio.data(), 8 * ratio.size());
    return std::fclose(f) == 0 && ok;
}

// Fills `ps` (ids = row index) and `time`; false on a missing or malformed file.
inline bool read(const std::filesystem::path& path, double& time, std::vector<Particle>& ps, double density) {
    std::FILE* f = std::fopen(path.string().c_str(), "rb");
    if (!f) return false;
    char magic[8];
    std::int64_t n = 0;
    bool ok = std::fread(magic, 1, 8, f) == 8 && std::memcmp(magic, "MPSBST1", 8) == 0 && std::fread(&n, sizeof n, 1, f) == 1 &&
              std::fread(&time, sizeof time, 1, f) == 1 && n >= 0 && n <= 0x7fffffff;
    if (!ok) {
        std::fclose(f);
        return false;
    }
    std::vector<std::int32_t> type((std::size_t) n), bc((std::size_t) n);
    std::vector<double> pos(3 * (std::size_t) n), vel(3 * (std::size_t) n), p((std::size_t) n), ratio((std::size_t) n);
    auto get = [&](void* d, std::size_t bytes) { ok = ok && (bytes == 0 || std::fread(d, 1, bytes, f) == bytes); };
    get(type.data(), 4 * type.size());
    get(bc.data(), 4 * bc.size());
    get(pos.data(), 8 * pos.size());
    get(vel.data(), 8 * vel.size());
    get(p.data(), 8 * p.size());
    get(ratio.data(), 8 * ratio.size());
    std::fclose(f);
    if (!ok) return false;
    ps.clear();
    ps.reserve((std::size_t) n);
    for (std::size_t i = 0; i < (std::size_t) n; i++) {
        ps.push_back(Particle((int) i, static_cast<ParticleType>(type[i]), Eigen::Vector3d(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]),
                              Eigen::Vector3d(vel[3 * i], vel[3 * i + 1], vel[3 * i + 2]), density));
        ps.back().pressure           = p[i];
        ps.back().numberDensityRatio = ratio[i];
        ps.back().boundaryCondition  = static_cast<BoundaryCondition>(bc[i]);
    }
    return true;
}

} // namespace state_io
