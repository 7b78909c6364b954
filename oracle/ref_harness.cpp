// ref_harness.cpp — TEST INFRASTRUCTURE, NOT PRODUCT.
//
// Thin C-ABI driver around the UNMODIFIED reference classes, compiled from the sources where
// they lie under $MPS_REF_DIR (default /root/reference) into oracle/_ref/libmps_ref.so by
// oracle/Makefile. The only "patch" is made on a temporary copy of src/mps.hpp in the build
// directory (`private:` -> `public:`) so the phases of MPS::stepForward (src/mps.cpp:111-130)
// can be called one at a time. Nothing here is imported by the product path.
#include "mps.hpp" // the sed-patched temporary copy (oracle/_ref/build/patched/mps.hpp)

#include "../include/mps_b200.h"

#include <Eigen/IterativeLinearSolvers>
#include <cstring>
#include <filesystem>

namespace {
struct Ref {
    MPS mps;
    double initialTime = 0;
};
} // namespace

extern "C" {

// MPS(yaml) + loadInitialState (src/mps.cpp:18-36, 96-109). `csv_path` overrides
// settings.inputCsvPath (which the reference resolves against the cwd, SURVEY B-12).
void* ref_create(const char* yaml_path, const char* csv_path) {
    auto* r = new Ref{MPS(std::filesystem::path(yaml_path))};
    if (csv_path && csv_path[0]) r->mps.settings.inputCsvPath = csv_path;
    r->initialTime = r->mps.loadInitialState();
    // SURVEY B-1: pressure / boundaryCondition are uninitialised in the reference; define them.
    for (auto& p : r->mps.particles) {
        p.pressure = 0.0;
    }
    return r;
}

// Same, but particles come from arrays (for generated cases that a 6-decimal CSV cannot carry).
void* ref_create_from_arrays(const char* yaml_path, int64_t n, const int32_t* type, const double* pos,
                             const double* vel) {
    auto* r = new Ref{MPS(std::filesystem::path(yaml_path))};
    MPS& m  = r->mps;
    for (int64_t i = 0; i < n; i++) {
        Particle p((int) i, static_cast<ParticleType>(type[i]),
                   Eigen::Vector3d(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]),
                   Eigen::Vector3d(vel[3 * i], vel[3 * i + 1], vel[3 * i + 2]), m.settings.density);
        p.pressure             = 0.0;
        p.boundaryCondition    = BoundaryCondition::Ignored;
        p.isDirichletBoundaryConnected = false;
        m.particles.push_back(p);
    }
    // the remainder of loadInitialState (src/mps.cpp:99-106)
    m.sourceTerm.resize(m.particles.size());
    m.coefficientMatrix.resize(m.particles.size(), m.particles.size());
    m.bucket = Bucket(m.settings.effectiveRadius.max, m.settings.domain, m.particles.size());
    m.setNumberDensity();
    return r;
}

void ref_destroy(void* h) { delete static_cast<Ref*>(h); }
int64_t ref_num_particles(void* h) { return (int64_t) static_cast<Ref*>(h)->mps.particles.size(); }
double ref_initial_time(void* h) { return static_cast<Ref*>(h)->initialTime; }
int ref_debug_log_count(void* h) { return static_cast<Ref*>(h)->mps.debugLogCount; }

void ref_get_settings(void* h, mpsb_settings* s) {
    MPS& m = static_cast<Ref*>(h)->mps;
    const Settings& t = m.settings;
    std::memset(s, 0, sizeof *s);
    s->dim                      = (int) t.dim;
    s->quasi_compressibility_on = t.quasiCompressibility.on;
    s->higher_order_source_on   = t.higherOrderSourceTerm.on;
    s->particle_distance        = t.particleDistance;
    s->dt                       = t.dt;
    s->cfl_condition            = t.cflCondition;
    s->domain_min[0] = t.domain.x.min; s->domain_max[0] = t.domain.x.max;
    s->domain_min[1] = t.domain.y.min; s->domain_max[1] = t.domain.y.max;
    s->domain_min[2] = t.domain.z.min; s->domain_max[2] = t.domain.z.max;
    s->density             = t.density;
    s->kinematic_viscosity = t.kinematicViscosity;
    for (int k = 0; k < 3; k++) s->gravity[k] = t.gravity[k];
    s->re_pressure       = t.effectiveRadius.pressure;
    s->re_viscosity      = t.effectiveRadius.viscosity;
    s->re_surface        = t.effectiveRadius.surfaceDetection;
    s->re_max            = t.effectiveRadius.max;
    s->surface_threshold = t.thresholdForSurfaceDetection;
    s->compressibility   = t.quasiCompressibility.on ? t.quasiCompressibility.compressibility : 0.0;
    s->higher_order_gamma = t.higherOrderSourceTerm.on ? t.higherOrderSourceTerm.gamma : 0.0;
    s->relaxation_coefficient = t.relaxationCoefficientForPressure;
    s->collision_distance     = t.collisionDistance;
    s->restitution            = t.coefficientOfRestitution;
    s->n0_pressure      = m.refValues.pressure.initialNumberDensity;
    s->lambda_pressure  = m.refValues.pressure.lambda;
    s->n0_viscosity     = m.refValues.viscosity.initialNumberDensity;
    s->lambda_viscosity = m.refValues.viscosity.lambda;
    s->n0_surface       = m.refValues.surfaceDetection.initialNumberDensity;
    s->lambda_surface   = m.refValues.surfaceDetection.lambda;
}

// Override a few settings after load (variant configs without writing new yml files).
void ref_set_pressure_options(void* h, int qc_on, double alpha, int ho_on, double gamma, double relax) {
    Settings& t = static_cast<Ref*>(h)->mps.settings;
    t.quasiCompressibility.on              = qc_on;
    t.quasiCompressibility.compressibility = alpha;
    t.higherOrderSourceTerm.on             = ho_on;
    t.higherOrderSourceTerm.gamma          = gamma;
    t.relaxationCoefficientForPressure     = relax;
}

void ref_get_state(void* h, const mpsb_fields* f) {
    MPS& m = static_cast<Ref*>(h)->mps;
    size_t n = m.particles.size();
    for (size_t i = 0; i < n; i++) {
        const Particle& p = m.particles[i];
        if (f->type) f->type[i] = (int) p.type;
        for (int k = 0; k < 3; k++) {
            if (f->position) f->position[3 * i + k] = p.position[k];
            if (f->velocity) f->velocity[3 * i + k] = p.velocity[k];
            if (f->acceleration) f->acceleration[3 * i + k] = p.acceleration[k];
        }
        if (f->pressure) f->pressure[i] = p.pressure;
        if (f->minimum_pressure) f->minimum_pressure[i] = p.minimumPressure;
        if (f->number_density_ratio) f->number_density_ratio[i] = p.numberDensityRatio;
        if (f->boundary_condition) f->boundary_condition[i] = (int) p.boundaryCondition;
    }
}

void ref_set_state(void* h, const mpsb_fields* f) {
    MPS& m = static_cast<Ref*>(h)->mps;
    size_t n = m.particles.size();
    for (size_t i = 0; i < n; i++) {
        Particle& p = m.particles[i];
        if (f->type) p.type = static_cast<ParticleType>(f->type[i]);
        for (int k = 0; k < 3; k++) {
            if (f->position) p.position[k] = f->position[3 * i + k];
            if (f->velocity) p.velocity[k] = f->velocity[3 * i + k];
            if (f->acceleration) p.acceleration[k] = f->acceleration[3 * i + k];
        }
        if (f->pressure) p.pressure = f->pressure[i];
        if (f->minimum_pressure) p.minimumPressure = f->minimum_pressure[i];
        if (f->number_density_ratio) p.numberDensityRatio = f->number_density_ratio[i];
        if (f->boundary_condition) p.boundaryCondition = static_cast<BoundaryCondition>(f->boundary_condition[i]);
    }
}

// One private phase of the reference; ids follow mpsb_phase where they coincide, finer where not.
//  0 setNeighbors  1 calcGravity  2 calcViscosity  3 moveParticles  4 collision
//  5 setBoundaryCondition  6 setSourceTerm  7 setMatrix  8 ensureDirichletBoundaryConnection
//  9 solvePoissonEquation  10 removeNegativePressure  11 setMinimumPressure
//  12 calcPressureGradient  13 moveParticlesWithPressureGradient  14 setNumberDensity
//  15 getCourantNumber (returned)
double ref_phase(void* h, int phase) {
    MPS& m = static_cast<Ref*>(h)->mps;
    switch (phase) {
    case 0: m.setNeighbors(); break;
    case 1: m.calcGravity(); break;
    case 2: m.calcViscosity(); break;
    case 3: m.moveParticles(); break;
    case 4: m.collision(); break;
    case 5: m.setBoundaryCondition(); break;
    case 6: m.setSourceTerm(); break;
    case 7: m.setMatrix(); break;
    case 8: m.ensureDirichletBoundaryConnection(); break;
    case 9: m.solvePoissonEquation(); break;
    case 10: m.removeNegativePressure(); break;
    case 11: m.setMinimumPressure(); break;
    case 12: m.calcPressureGradient(); break;
    case 13: m.moveParticlesWithPressureGradient(); break;
    case 14: m.setNumberDensity(); break;
    case 15: return m.getCourantNumber();
    default: return -1.0;
    }
    return 0.0;
}

double ref_step(void* h, int is_time_to_export) {
    return static_cast<Ref*>(h)->mps.stepForward(is_time_to_export != 0);
}

// Neighbour lists as CSR by id, in the reference's stored order.
int64_t ref_get_neighbors(void* h, int64_t* row_start, int32_t* ids, double* dist, int64_t capacity) {
    MPS& m = static_cast<Ref*>(h)->mps;
    int64_t k = 0;
    size_t n  = m.particles.size();
    for (size_t i = 0; i < n; i++) {
        if (row_start) row_start[i] = k;
        // a Ghost keeps its stale list (src/mps.cpp:590-593); report it as empty
        if (m.particles[i].type == ParticleType::Ghost) continue;
        for (auto& nb : m.particles[i].neighbors) {
            if (ids && k < capacity) {
                ids[k]  = nb.id;
                dist[k] = nb.distance;
            }
            k++;
        }
    }
    if (row_start) row_start[n] = k;
    return k;
}

// coefficientMatrix (RowMajor CSR after setFromTriplets, src/mps.cpp:359) and sourceTerm.
int64_t ref_get_ppe(void* h, int64_t* row_start, int32_t* col, double* val, double* rhs, int64_t capacity) {
    MPS& m  = static_cast<Ref*>(h)->mps;
    auto& A = m.coefficientMatrix;
    int64_t k = 0;
    for (int i = 0; i < A.outerSize(); i++) {
        if (row_start) row_start[i] = k;
        for (Eigen::SparseMatrix<double, Eigen::RowMajor>::InnerIterator it(A, i); it; ++it) {
            if (col && k < capacity) {
                col[k] = it.col();
                val[k] = it.value();
            }
            k++;
        }
    }
    if (row_start) row_start[A.outerSize()] = k;
    if (rhs)
        for (int i = 0; i < m.sourceTerm.size(); i++) rhs[i] = m.sourceTerm[i];
    return k;
}

// The same solver object the reference constructs (src/mps.cpp:452-457) on the current
// coefficientMatrix/sourceTerm, so iterations()/error() can be read. x may be NULL.
int ref_cg_solve(void* h, double* x, double* error, int* info) {
    MPS& m = static_cast<Ref*>(h)->mps;
    Eigen::ConjugateGradient<Eigen::SparseMatrix<double>> solver;
    solver.compute(m.coefficientMatrix);
    Eigen::VectorXd p = solver.solve(m.sourceTerm);
    if (x)
        for (int i = 0; i < p.size(); i++) x[i] = p[i];
    if (error) *error = solver.error();
    if (info) *info = (int) solver.info();
    return (int) solver.iterations();
}

void ref_get_dirichlet_connected(void* h, int32_t* out) {
    MPS& m = static_cast<Ref*>(h)->mps;
    for (size_t i = 0; i < m.particles.size(); i++) out[i] = m.particles[i].isDirichletBoundaryConnected ? 1 : 0;
}

void ref_bucket_dims(void* h, int32_t* nxyz) {
    MPS& m  = static_cast<Ref*>(h)->mps;
    nxyz[0] = m.bucket.numX;
    nxyz[1] = m.bucket.numY;
    nxyz[2] = m.bucket.numZ;
}

} // extern "C"
