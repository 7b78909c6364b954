#include "mps.hpp"

#include "../../include/mps_b200.h"
#include "csv_io.hpp"
#include "log.hpp"
#include "rendezvous.hpp"
#include "state_io.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>

namespace {
// reference weight(), src/mps.cpp:9-16
inline double weightOf(double dist, double re) { return dist < re ? (re / dist) - 1.0 : 0.0; }

int columnIndex(const std::vector<std::string>& header, const std::string& name) {
    for (std::size_t i = 0; i < header.size(); i++)
        if (header[i] == name) return (int) i;
    return -1;
}
} // namespace

RefValues::RefValues(const int& dim, const double& particleDistance, const double& effectiveRadius) {
    const int zLo = dim == 2 ? 0 : -4, zHi = dim == 2 ? 1 : 5;
    for (int iX = -4; iX < 5; iX++)
        for (int iY = -4; iY < 5; iY++)
            for (int iZ = zLo; iZ < zHi; iZ++) {
                if (iX == 0 && iY == 0 && iZ == 0) continue;
                const Eigen::Vector3d r(particleDistance * (double) iX, particleDistance * (double) iY, particleDistance * (double) iZ);
                const double dist  = r.norm(); // sqrt((x*x + y*y) + z*z)
                const double dist2 = dist * dist;
                const double w     = weightOf(dist, effectiveRadius);
                initialNumberDensity += w;
                lambda += dist2 * w;
            }
    lambda /= initialNumberDensity;
}

MPS::MPS(const std::filesystem::path& inputYamlPath) {
    settings.load(inputYamlPath);
    refValues.pressure         = RefValues((int) settings.dim, settings.particleDistance, settings.effectiveRadius.pressure);
    refValues.viscosity        = RefValues((int) settings.dim, settings.particleDistance, settings.effectiveRadius.viscosity);
    refValues.surfaceDetection = RefValues((int) settings.dim, settings.particleDistance, settings.effectiveRadius.surfaceDetection);
}

void MPS::fail(const char* what, int status) {
    mpslog::error(std::string(what) + " failed (status " + std::to_string(status) + "): " + mpsb_last_error(handle.get()));
    mpslog::error("The MPS step has no CPU path; a CUDA device (B200, sm_100a) is required.");
    std::exit(EXIT_FAILURE);
}

double MPS::importInitialCondition() {
    if (!std::filesystem::exists(settings.inputCsvPath)) {
        mpslog::error("Input file does not exist in the specified path: " + settings.inputCsvPath.string());
        std::exit(EXIT_FAILURE);
    }
    std::ifstream in(settings.inputCsvPath);
    std::string line;
    double initialTime = 0.0;
    std::vector<std::string> header;
    for (int rowNumber = 1; std::getline(in, line); rowNumber++) {
        auto f = csv_io::splitRecord(line);
        if (rowNumber == 1 && f.size() > 1) initialTime = std::strtod(f[1].c_str(), nullptr);
        if (rowNumber == 3) {
            header = f;
            break;
        }
    }
    const int cId = columnIndex(header, "ID"), cType = columnIndex(header, "Type");
    const int cX = columnIndex(header, "Position.x (m)"), cY = columnIndex(header, "Position.y (m)"), cZ = columnIndex(header, "Position.z (m)");
    const int cU = columnIndex(header, "Velocity.x (m/s)"), cV = columnIndex(header, "Velocity.y (m/s)"), cW = columnIndex(header, "Velocity.z (m/s)");
    if (std::min({cId, cType, cX, cY, cZ, cU, cV, cW}) < 0) {
        mpslog::error("Input file lacks a particle column (ID, Type, Position.*, Velocity.*): " + settings.inputCsvPath.string());
        std::exit(EXIT_FAILURE);
    }
    particles.clear();
    while (std::getline(in, line)) {
        if (line.empty() || line == "\r") continue;
        auto f   = csv_io::splitRecord(line);
        auto num = [&](int c) { return std::strtod(f.at((std::size_t) c).c_str(), nullptr); };
        particles.push_back(Particle(
            (int) std::strtol(f.at((std::size_t) cId).c_str(), nullptr, 10),
            static_cast<ParticleType>((int) std::strtol(f.at((std::size_t) cType).c_str(), nullptr, 10)),
            Eigen::Vector3d(num(cX), num(cY), num(cZ)), Eigen::Vector3d(num(cU), num(cV), num(cW)), settings.density
        ));
    }
    return initialTime;
}

void MPS::createDeviceState(bool restoreDerived) {
    mpsb_settings s;
    std::memset(&s, 0, sizeof s);
    s.dim                      = (int) settings.dim;
    s.quasi_compressibility_on = settings.quasiCompressibility.on ? 1 : 0;
    s.higher_order_source_on   = settings.higherOrderSourceTerm.on ? 1 : 0;
    s.particle_distance        = settings.particleDistance;
    s.dt                       = settings.dt;
    s.cfl_condition            = settings.cflCondition;
    s.domain_min[0] = settings.domain.x.min; s.domain_max[0] = settings.domain.x.max;
    s.domain_min[1] = settings.domain.y.min; s.domain_max[1] = settings.domain.y.max;
    s.domain_min[2] = settings.domain.z.min; s.domain_max[2] = settings.domain.z.max;
    s.density             = settings.density;
    s.kinematic_viscosity = settings.kinematicViscosity;
    for (int k = 0; k < 3; k++) s.gravity[k] = settings.gravity[(std::size_t) k];
    s.re_pressure            = settings.effectiveRadius.pressure;
    s.re_viscosity           = settings.effectiveRadius.viscosity;
    s.re_surface             = settings.effectiveRadius.surfaceDetection;
    s.re_max                 = settings.effectiveRadius.max;
    s.surface_threshold      = settings.thresholdForSurfaceDetection;
    s.compressibility        = settings.quasiCompressibility.on ? settings.quasiCompressibility.compressibility : 0.0;
    s.higher_order_gamma     = settings.higherOrderSourceTerm.on ? settings.higherOrderSourceTerm.gamma : 0.0;
    s.relaxation_coefficient = settings.relaxationCoefficientForPressure;
    s.collision_distance     = settings.collisionDistance;
    s.restitution            = settings.coefficientOfRestitution;
    s.n0_pressure  = refValues.pressure.initialNumberDensity;         s.lambda_pressure  = refValues.pressure.lambda;
    s.n0_viscosity = refValues.viscosity.initialNumberDensity;        s.lambda_viscosity = refValues.viscosity.lambda;
    s.n0_surface   = refValues.surfaceDetection.initialNumberDensity; s.lambda_surface   = refValues.surfaceDetection.lambda;

    const std::size_t n = particles.size();
    std::vector<std::int32_t> type(n);
    std::vector<double> pos(3 * n), vel(3 * n);
    for (std::size_t i = 0; i < n; i++) {
        const Particle& pi = particles[i];
        if (pi.id != (int) i) { // the reference indexes particles[neighbor.id] (src/mps.cpp:599-617)
            mpslog::error("Particle ids must equal their row index (row " + std::to_string(i) + " has id " + std::to_string(pi.id) + ").");
            std::exit(EXIT_FAILURE);
        }
        type[i] = static_cast<std::int32_t>(pi.type);
        for (std::size_t k = 0; k < 3; k++) {
            pos[3 * i + k] = pi.position[k];
            vel[3 * i + k] = pi.velocity[k];
        }
    }
    const rendezvous::Env env = rendezvous::fromEnvironment();
    rank      = env.rank;
    worldSize = env.world;
    if (worldSize > 1) device = env.localRank;
    if (const char* d = std::getenv("MPSB_DEVICE")) device = std::atoi(d);
    mpsb_handle* h = nullptr;
    int rc         = mpsb_create(&s, (std::int64_t) n, type.data(), pos.data(), vel.data(), device, &h);
    if (rc != MPSB_OK) {
        mpslog::error(std::string("mpsb_create failed (status ") + std::to_string(rc) + "): " + mpsb_last_error(nullptr));
        mpslog::error("The MPS step has no CPU path; a CUDA device (B200, sm_100a) is required.");
        std::exit(EXIT_FAILURE);
    }
    handle = std::shared_ptr<mpsb_handle>(h, [](mpsb_handle* p) { mpsb_destroy(p); });
    if (worldSize > 1) {
        unsigned char id[128] = {0};
        if (rank == 0 && mpsb_comm_unique_id(id) != MPSB_OK) {
            mpslog::error("mpsb_comm_unique_id failed: NCCL (libnccl.so.2) is required for multi-GPU runs");
            std::exit(EXIT_FAILURE);
        }
        if (!rendezvous::shareUniqueId(env, id)) {
            mpslog::error("Multi-GPU rendezvous failed (MASTER_ADDR=" + env.addr + ", port " + std::to_string(env.port + 1) + ")");
            std::exit(EXIT_FAILURE);
        }
        rc = mpsb_attach_comm(h, rank, worldSize, id);
        if (rc != MPSB_OK) fail("mpsb_attach_comm", rc);
    }
    // setNumberDensity at the beginning, as reference loadInitialState (src/mps.cpp:104-106)
    rc = mpsb_run_phase(h, MPSB_PH_NUMBER_DENSITY, nullptr);
    if (rc != MPSB_OK) fail("initial setNumberDensity", rc);
    if (restoreDerived) { // checkpoint: put back what the interrupted run held instead of the freshly derived values
        std::vector<std::int32_t> bc(n);
        std::vector<double> p(n), ratio(n);
        for (std::size_t i = 0; i < n; i++) {
            bc[i]    = static_cast<std::int32_t>(particles[i].boundaryCondition);
            p[i]     = particles[i].pressure;
            ratio[i] = particles[i].numberDensityRatio;
        }
        mpsb_fields f;
        std::memset(&f, 0, sizeof f);
        f.pressure = p.data(); f.number_density_ratio = ratio.data(); f.boundary_condition = bc.data();
        rc = mpsb_upload(h, MPSB_F_PRESSURE | MPSB_F_NUMBER_DENSITY_RATIO | MPSB_F_BOUNDARY_CONDITION, &f);
        if (rc != MPSB_OK) fail("mpsb_upload (checkpoint)", rc);
    }
    downloadParticles();
}

double MPS::loadInitialStateFromFile(const std::filesystem::path& statePath) {
    double time = 0.0;
    if (!state_io::read(statePath, time, particles, settings.density)) {
        mpslog::error("Could not read the state file: " + statePath.string());
        std::exit(EXIT_FAILURE);
    }
    createDeviceState(true);
    return time;
}

bool MPS::saveState(const std::filesystem::path& statePath, double time, bool download) {
    if (download) downloadParticles();
    return state_io::write(statePath, time, particles);
}

double MPS::loadInitialState() {
    double initialTime = importInitialCondition();
    createDeviceState();
    return initialTime;
}

double MPS::loadInitialStateFromArrays(
    double time, std::int64_t n, const std::int32_t* type, const double* position3, const double* velocity3
) {
    particles.clear();
    particles.reserve((std::size_t) n);
    for (std::int64_t i = 0; i < n; i++)
        particles.push_back(Particle(
            (int) i, static_cast<ParticleType>(type[i]), Eigen::Vector3d(position3[3 * i], position3[3 * i + 1], position3[3 * i + 2]),
            Eigen::Vector3d(velocity3[3 * i], velocity3[3 * i + 1], velocity3[3 * i + 2]), settings.density
        ));
    createDeviceState();
    return time;
}

void MPS::reportEvents(const mpsb_step_stats& st) {
    // texts of reference src/mps.cpp:655-662, :405-409, :560-568 (std::to_string spelling)
    if (st.n_ghosted > 0) {
        std::vector<std::int32_t> ids((std::size_t) st.n_ghosted);
        std::vector<double> xyz(3 * (std::size_t) st.n_ghosted);
        int count = 0;
        if (mpsb_events(handle.get(), 0, ids.data(), xyz.data(), st.n_ghosted, &count) == MPSB_OK)
            for (int k = 0; k < std::min(count, st.n_ghosted); k++)
                mpslog::debug(
                    "Particle (id=" + std::to_string(ids[(std::size_t) k]) + ") is out of domain. (x, y, z)=(" + std::to_string(xyz[3 * (std::size_t) k]) + ", " +
                    std::to_string(xyz[3 * (std::size_t) k + 1]) + ", " + std::to_string(xyz[3 * (std::size_t) k + 2]) + ") "
                );
    }
    if (st.n_unconnected > 0) {
        std::vector<std::int32_t> ids((std::size_t) st.n_unconnected);
        int count = 0;
        if (mpsb_events(handle.get(), 1, ids.data(), nullptr, st.n_unconnected, &count) == MPSB_OK)
            for (int k = 0; k < std::min(count, st.n_unconnected); k++)
                mpslog::debug("There is no Dirichlet boundary condition connected to the particle (id=" + std::to_string(ids[(std::size_t) k]) + ").");
    }
    if (st.courant_exceeds_cfl) mpslog::debug("Courant number is larger than CFL condition. Courant=" + std::to_string(st.courant));
    if (st.courant > 1.0) mpslog::error("Courant number is larger than 1.0. Exiting Simulation."); // the reference does not exit either (SURVEY B-5)
    debugLogCount += st.n_ghosted + st.n_unconnected + (st.courant_exceeds_cfl ? 1 : 0);
}

double MPS::stepForward(const bool isTimeToExport) {
    if (!handle) {
        mpslog::error("MPS::stepForward called before loadInitialState.");
        std::exit(EXIT_FAILURE);
    }
    mpsb_step_stats st;
    // setNumberDensity only when its result is consumed (reference src/mps.cpp:125-127)
    const int wantNumberDensity = (isTimeToExport || settings.higherOrderSourceTerm.on) ? 1 : 0;
    int rc = mpsb_step(handle.get(), wantNumberDensity, &st);
    if (rc != MPSB_OK) fail("mpsb_step", rc);
    reportEvents(st);
    lastStep.cgIterations  = st.cg_iterations;
    lastStep.nInner        = st.n_inner;
    lastStep.nSurface      = st.n_surface;
    lastStep.nnz           = st.nnz;
    lastStep.nActive       = st.n_active;
    lastStep.nCollisions   = st.n_collisions;
    lastStep.cgRelResidual = st.cg_rel_residual;
    lastStep.msTotal       = st.ms_total;
    lastStep.msSolve       = st.ms_solve;
    lastStep.msNeighbors   = st.ms_neighbors;
    if (isTimeToExport || syncEveryStep) downloadParticles();
    return st.courant;
}

void MPS::downloadParticles() {
    if (!handle) return;
    const std::size_t n = particles.size();
    // one page-locked block, reused by every export (f-1: pinned D2H instead of pageable vectors): 12 doubles + 2 ints per particle
    const std::size_t need = n * (12 * sizeof(double) + 2 * sizeof(std::int32_t)) + 64;
    if (stagingBytes < need) {
        staging.reset();
        void* block = mpsb_host_alloc(need);
        if (!block) fail("mpsb_host_alloc (pinned staging)", MPSB_ERR_CUDA);
        staging      = std::shared_ptr<void>(block, [](void* p) { mpsb_host_free(p); });
        stagingBytes = need;
    }
    double* d = static_cast<double*>(staging.get());
    double *pos = d, *vel = d + 3 * n, *acc = d + 6 * n, *p = d + 9 * n, *minp = d + 10 * n, *ratio = d + 11 * n;
    std::int32_t *type = reinterpret_cast<std::int32_t*>(d + 12 * n), *bc = type + n;
    mpsb_fields f;
    f.type = type; f.position = pos; f.velocity = vel; f.acceleration = acc;
    f.pressure = p; f.minimum_pressure = minp; f.number_density_ratio = ratio; f.boundary_condition = bc;
    int rc = mpsb_download(handle.get(), MPSB_F_ALL, &f);
    if (rc != MPSB_OK) fail("mpsb_download", rc);
    for (std::size_t i = 0; i < n; i++) {
        Particle& pi = particles[i];
        pi.type      = static_cast<ParticleType>(type[i]);
        for (std::size_t k = 0; k < 3; k++) {
            pi.position[k]     = pos[3 * i + k];
            pi.velocity[k]     = vel[3 * i + k];
            pi.acceleration[k] = acc[3 * i + k];
        }
        pi.pressure           = p[i];
        pi.minimumPressure    = minp[i];
        pi.numberDensityRatio = ratio[i];
        pi.boundaryCondition  = static_cast<BoundaryCondition>(bc[i]);
    }
}

void MPS::uploadParticles() {
    if (!handle) return;
    const std::size_t n = particles.size();
    std::vector<std::int32_t> type(n);
    std::vector<double> pos(3 * n), vel(3 * n);
    for (std::size_t i = 0; i < n; i++) {
        type[i] = static_cast<std::int32_t>(particles[i].type);
        for (std::size_t k = 0; k < 3; k++) {
            pos[3 * i + k] = particles[i].position[k];
            vel[3 * i + k] = particles[i].velocity[k];
        }
    }
    mpsb_fields f;
    std::memset(&f, 0, sizeof f);
    f.type = type.data(); f.position = pos.data(); f.velocity = vel.data();
    int rc = mpsb_upload(handle.get(), MPSB_F_TYPE | MPSB_F_POSITION | MPSB_F_VELOCITY, &f);
    if (rc != MPSB_OK) fail("mpsb_upload", rc);
}

void MPS::downloadNeighbors() {
    if (!handle) return;
    const std::size_t n = particles.size();
    std::vector<std::int64_t> rowStart(n + 1);
    std::int64_t total = 0;
    int rc = mpsb_get_neighbors(handle.get(), rowStart.data(), nullptr, nullptr, 0, &total);
    if (rc != MPSB_OK) fail("mpsb_get_neighbors", rc);
    std::vector<std::int32_t> ids((std::size_t) total);
    std::vector<double> dist((std::size_t) total);
    rc = mpsb_get_neighbors(handle.get(), rowStart.data(), ids.data(), dist.data(), total, &total);
    if (rc != MPSB_OK) fail("mpsb_get_neighbors", rc);
    for (std::size_t i = 0; i < n; i++) {
        particles[i].neighbors.clear();
        for (std::int64_t k = rowStart[i]; k < rowStart[i + 1]; k++) particles[i].neighbors.emplace_back(ids[(std::size_t) k], dist[(std::size_t) k]);
    }
}
