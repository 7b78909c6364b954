// host_selftest.cpp — CPU-only driver for the host code (no device calls), used by tests/test_host.py:
//   host_selftest settings <settings.yml>            parsed Settings + RefValues, one "key value" per line (%.17g)
//   host_selftest reexport <settings.yml> <out dir>  input CSV -> particles -> out/particles.csv, out/particles.vtu
//   host_selftest float <style> <value>...           spelling of doubles (style: fixed6 | gcc5)
//   host_selftest rendezvous x                       RANK/WORLD_SIZE/MASTER_* from the environment: share a 128-byte id, print its sum
//   host_selftest exportbench <n> <out dir>          time Exporter::toCsv / toVtu on n synthetic particles (SURVEY f-1)
//   host_selftest state <n> <file>                   binary state: write n particles, read back, compare bit for bit;
//                                                    a truncated copy and a wrong magic must be refused
#include "exporter.hpp"
#include "mps.hpp"
#include "rendezvous.hpp"
#include "state_io.hpp"

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

static void kv(const char* k, double v) { std::printf("%s %.17g\n", k, v); }

int main(int argc, char** argv) {
    if (argc < 3) {
        std::fprintf(stderr, "usage: host_selftest settings|reexport|float ...\n");
        return 2;
    }
    const std::string mode = argv[1];
    if (mode == "float") {
        auto style = std::strcmp(argv[2], "gcc5") == 0 ? csv_io::FloatStyle::Trunc5Gcc : csv_io::FloatStyle::Fixed6;
        for (int i = 3; i < argc; i++) std::printf("%s\n", csv_io::formatFloat(std::strtod(argv[i], nullptr), style).c_str());
        return 0;
    }
    if (mode == "rendezvous") {
        const rendezvous::Env env = rendezvous::fromEnvironment();
        unsigned char id[128] = {0};
        if (env.rank == 0)
            for (int i = 0; i < 128; i++) id[i] = (unsigned char) (i * 7 + 3);
        if (!rendezvous::shareUniqueId(env, id)) {
            std::printf("rank %d rendezvous failed\n", env.rank);
            return 1;
        }
        unsigned sum = 0;
        for (int i = 0; i < 128; i++) sum += id[i] * (unsigned) (i + 1);
        std::printf("rank %d of %d sum %u\n", env.rank, env.world, sum);
        return 0;
    }
    if (mode == "state" && argc >= 4) {
        const long n = std::atol(argv[2]);
        const std::filesystem::path path(argv[3]);
        std::vector<Particle> ps;
        for (long i = 0; i < n; i++) {
            // awkward values on purpose: denormals, negative zero, values no decimal CSV would carry
            const double x = 1e-4 * (double) i / 3.0, y = -0.0, z = 4.9e-324 * (double) (i % 3);
            ps.push_back(Particle((int) i, static_cast<ParticleType>((int) (i % 4) - 1), Eigen::Vector3d(x, y, z),
                                  Eigen::Vector3d(-x, 1.0 / (double) (i + 1), 0.1 + (double) i), 1000.0));
            ps.back().pressure           = 1e300 / (double) (i + 1);
            ps.back().numberDensityRatio = 1.0 - 1e-16 * (double) i;
            ps.back().boundaryCondition  = static_cast<BoundaryCondition>((int) (i % 3));
        }
        const double time = 0.1 + 0.2; // 0.30000000000000004
        if (!state_io::write(path, time, ps)) return 1;
        std::vector<Particle> back;
        double t2 = 0.0;
        if (!state_io::read(path, t2, back, 1000.0) || back.size() != ps.size() || std::memcmp(&t2, &time, 8) != 0) return 1;
        long bad = 0;
        for (std::size_t i = 0; i < ps.size(); i++) {
            const Particle &a = ps[i], &b = back[i];
            bool same = a.id == b.id && a.type == b.type && a.boundaryCondition == b.boundaryCondition &&
                        std::memcmp(&a.pressure, &b.pressure, 8) == 0 &&
                        std::memcmp(&a.numberDensityRatio, &b.numberDensityRatio, 8) == 0;
            for (int k = 0; k < 3; k++) {
                const double pa = a.position[k], pb = b.position[k], va = a.velocity[k], vb = b.velocity[k];
                same = same && std::memcmp(&pa, &pb, 8) == 0 && std::memcmp(&va, &vb, 8) == 0;
            }
            bad += !same;
        }
        // refusals: a file cut short, and a file that is not a state file
        const auto bytes = std::filesystem::file_size(path);
        std::filesystem::path cut = path;
        cut += ".cut";
        std::filesystem::copy_file(path, cut, std::filesystem::copy_options::overwrite_existing);
        if (n > 0) std::filesystem::resize_file(cut, bytes - 5);
        const bool cut_refused = n == 0 || !state_io::read(cut, t2, back, 1000.0);
        std::FILE* f = std::fopen(cut.string().c_str(), "wb");
        std::fputs("id,type,x,y,z\n0,0,0,0,0\n", f);
        std::fclose(f);
        const bool text_refused = !state_io::read(cut, t2, back, 1000.0);
        std::filesystem::remove(cut);
        std::printf("particles %ld bytes %llu mismatches %ld truncated_refused %d text_refused %d\n", n,
                    (unsigned long long) bytes, bad, (int) cut_refused, (int) text_refused);
        return bad == 0 && cut_refused && text_refused ? 0 : 1;
    }
    if (mode == "exportbench" && argc >= 4) {
        const long n = std::atol(argv[2]);
        std::vector<Particle> ps;
        ps.reserve((std::size_t) n);
        for (long i = 0; i < n; i++) {
            const double x = 1e-4 * (double) (i % 2832), y = 1e-4 * (double) (i / 2832);
            ps.push_back(Particle((int) i, ParticleType::Fluid, Eigen::Vector3d(x, y, 0.0), Eigen::Vector3d(0.1 * x, -0.2 * y, 0.0), 1000.0));
            ps.back().pressure           = 9800.0 * (0.2 - y);
            ps.back().numberDensityRatio = 0.97 + 1e-9 * (double) i;
        }
        Exporter e;
        const std::filesystem::path dir(argv[3]);
        auto t0 = std::chrono::steady_clock::now();
        e.toCsv(dir / "bench.csv", ps, 1.25);
        auto t1 = std::chrono::steady_clock::now();
        e.toVtu(dir / "bench.vtu", ps, 1.25);
        auto t2 = std::chrono::steady_clock::now();
        const double csvS = std::chrono::duration<double>(t1 - t0).count(), vtuS = std::chrono::duration<double>(t2 - t1).count();
        std::printf("particles %ld csv_s %.3f csv_bytes %ju vtu_s %.3f vtu_bytes %ju\n", n, csvS,
                    (std::uintmax_t) std::filesystem::file_size(dir / "bench.csv"), vtuS,
                    (std::uintmax_t) std::filesystem::file_size(dir / "bench.vtu"));
        return 0;
    }
    MPS mps{std::filesystem::path(argv[2])};
    const Settings& s = mps.settings;
    if (mode == "settings") {
        kv("dim", s.dim); kv("particle_distance", s.particleDistance); kv("dt", s.dt); kv("finish_time", s.finishTime);
        kv("output_interval", s.outputInterval); kv("cfl_condition", s.cflCondition);
        kv("domain_min_x", s.domain.x.min); kv("domain_max_x", s.domain.x.max); kv("domain_min_y", s.domain.y.min);
        kv("domain_max_y", s.domain.y.max); kv("domain_min_z", s.domain.z.min); kv("domain_max_z", s.domain.z.max);
        kv("density", s.density); kv("kinematic_viscosity", s.kinematicViscosity);
        kv("gravity_x", s.gravity.x()); kv("gravity_y", s.gravity.y()); kv("gravity_z", s.gravity.z());
        kv("re_pressure", s.effectiveRadius.pressure); kv("re_viscosity", s.effectiveRadius.viscosity);
        kv("re_surface", s.effectiveRadius.surfaceDetection); kv("re_max", s.effectiveRadius.max);
        kv("surface_threshold", s.thresholdForSurfaceDetection);
        kv("quasi_compressibility_on", s.quasiCompressibility.on); kv("compressibility", s.quasiCompressibility.compressibility);
        kv("higher_order_source_on", s.higherOrderSourceTerm.on); kv("higher_order_gamma", s.higherOrderSourceTerm.gamma);
        kv("relaxation_coefficient", s.relaxationCoefficientForPressure);
        kv("collision_distance", s.collisionDistance); kv("restitution", s.coefficientOfRestitution);
        kv("n0_pressure", mps.refValuesPressure().initialNumberDensity); kv("lambda_pressure", mps.refValuesPressure().lambda);
        kv("n0_viscosity", mps.refValuesViscosity().initialNumberDensity); kv("lambda_viscosity", mps.refValuesViscosity().lambda);
        kv("n0_surface", mps.refValuesSurfaceDetection().initialNumberDensity); kv("lambda_surface", mps.refValuesSurfaceDetection().lambda);
        std::printf("input_csv_path %s\n", s.inputCsvPath.string().c_str());
        return 0;
    }
    if (mode == "reexport" && argc >= 4) {
        double t = mps.loadParticlesOnly();
        Exporter e;
        e.toCsv(std::filesystem::path(argv[3]) / "particles.csv", mps.particles, t);
        e.toVtu(std::filesystem::path(argv[3]) / "particles.vtu", mps.particles, t);
        std::printf("particles %zu time %.17g\n", mps.particles.size(), t);
        return 0;
    }
    return 2;
}
