// mps_oracle.cpp — TEST INFRASTRUCTURE, NOT PRODUCT. CPU restatement of the reference MPS step.
//
// Plain C++ (no Eigen), written from the arithmetic specification in SURVEY.md Appendix A; each
// function cites the reference lines it restates. Only tests/, bench.py's cpu_baseline leg and
// __graft_entry__.smoke() may load this library. Parity status: PINNED — tests/test_oracle.py
// checks every phase of this file bit-for-bit against oracle/_ref (the unmodified reference
// compiled from /root/reference) and against the golden dumps under tests/golden/ that were
// generated from it (tests/golden/make_golden.py).
//
// Build: g++ -std=c++17 -O2 -ffp-contract=off (no -march: SSE2 baseline, no FMA — SURVEY A.0).
#include "../include/mps_b200.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <queue>
#include <vector>

namespace {

struct Nb {
    int id;
    double dist;
};

struct V3 {
    double x, y, z;
};
inline V3 sub(const V3& a, const V3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
// Eigen Vector3d::dot / squaredNorm reduce as (c0 + c1) + c2
inline double dot(const V3& a, const V3& b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline double norm(const V3& a) { return std::sqrt(dot(a, a)); }

struct Oracle {
    mpsb_settings s;
    int n = 0;
    std::vector<int> type, bc, connected;
    std::vector<V3> pos, vel, acc;
    std::vector<double> pressure, minp, ratio;
    std::vector<std::vector<Nb>> nbrs;
    // bucket (src/bucket.hpp:8-19)
    int numX = 0, numY = 0, numZ = 0, num = 0;
    std::vector<int> first, last, next;
    // PPE: CSR sorted by column (Eigen setFromTriplets result) + rhs
    std::vector<int64_t> rowStart;
    std::vector<int> col;
    std::vector<double> val, rhs;
    int debugLogCount = 0;
    int cgIterations  = 0;
    double cgError    = 0;
};

// weight (src/mps.cpp:9-16)
inline double weight(double dist, double re) { return dist < re ? (re / dist) - 1.0 : 0.0; }

// Bucket::Bucket (src/bucket.cpp:10-22)
void bucketInit(Oracle& o) {
    double len = o.s.re_max;
    o.numX = (int) ((o.s.domain_max[0] - o.s.domain_min[0]) / len) + 3;
    o.numY = (int) ((o.s.domain_max[1] - o.s.domain_min[1]) / len) + 3;
    o.numZ = (int) ((o.s.domain_max[2] - o.s.domain_min[2]) / len) + 3;
    o.num  = o.numX * o.numY * o.numZ;
    o.first.assign(o.num, -1);
    o.last.assign(o.num, -1);
    o.next.assign(o.n, -1);
}

inline void cellOf(const Oracle& o, const V3& p, int& ix, int& iy, int& iz) {
    double len = o.s.re_max;
    ix = (int) ((p.x - o.s.domain_min[0]) / len) + 1;
    iy = (int) ((p.y - o.s.domain_min[1]) / len) + 1;
    iz = (int) ((p.z - o.s.domain_min[2]) / len) + 1;
}

// Bucket::storeParticles (src/bucket.cpp:24-51): linked lists in particle (id) order, Ghost skipped
void storeParticles(Oracle& o) {
    std::fill(o.first.begin(), o.first.end(), -1);
    std::fill(o.last.begin(), o.last.end(), -1);
    std::fill(o.next.begin(), o.next.end(), -1);
    for (int i = 0; i < o.n; i++) {
        if (o.type[i] == MPSB_GHOST) continue;
        int ix, iy, iz;
        cellOf(o, o.pos[i], ix, iy, iz);
        int b = ix + iy * o.numX + iz * o.numX * o.numY;
        if (o.last[b] == -1)
            o.first[b] = i;
        else
            o.next[o.last[b]] = i;
        o.last[b] = i;
    }
}

// MPS::setNeighbors (src/mps.cpp:585-619): jx outer, jy, jz inner; cell list in id order;
// a Ghost keeps its stale list (:590-593)
void setNeighbors(Oracle& o) {
    storeParticles(o);
    for (int i = 0; i < o.n; i++) {
        if (o.type[i] == MPSB_GHOST) continue;
        o.nbrs[i].clear();
        int ix, iy, iz;
        cellOf(o, o.pos[i], ix, iy, iz);
        for (int jx = ix - 1; jx <= ix + 1; jx++)
            for (int jy = iy - 1; jy <= iy + 1; jy++)
                for (int jz = iz - 1; jz <= iz + 1; jz++) {
                    int b = jx + jy * o.numX + jz * o.numX * o.numY;
                    for (int j = o.first[b]; j != -1; j = o.next[j]) {
                        double dist = norm(sub(o.pos[j], o.pos[i]));
                        if (j != i && dist < o.s.re_max) o.nbrs[i].push_back({j, dist});
                    }
                }
    }
}

// MPS::getNumberDensity (src/mps.cpp:573-583)
double numberDensity(const Oracle& o, int i, double re) {
    double n = 0.0;
    for (const Nb& nb : o.nbrs[i])
        if (nb.dist < re) n += weight(nb.dist, re);
    return n;
}

// MPS::setNumberDensity (src/mps.cpp:627-641)
void setNumberDensity(Oracle& o) {
    setNeighbors(o);
    for (int i = 0; i < o.n; i++)
        o.ratio[i] = o.bc[i] == MPSB_BC_SURFACE ? 0.0 : numberDensity(o, i, o.s.re_pressure) / o.s.n0_pressure;
}

// MPS::calcGravity (src/mps.cpp:165-175)
void calcGravity(Oracle& o) {
    for (int i = 0; i < o.n; i++)
        o.acc[i] = o.type[i] == MPSB_FLUID ? V3{o.s.gravity[0], o.s.gravity[1], o.s.gravity[2]} : V3{0, 0, 0};
}

// MPS::calcViscosity (src/mps.cpp:177-203)
void calcViscosity(Oracle& o) {
    const double re = o.s.re_viscosity;
    double a = (o.s.kinematic_viscosity) * (2.0 * (double) o.s.dim) / (o.s.n0_viscosity * o.s.lambda_viscosity);
    for (int i = 0; i < o.n; i++) {
        if (o.type[i] != MPSB_FLUID) continue;
        V3 t{0, 0, 0};
        for (const Nb& nb : o.nbrs[i]) {
            if (nb.dist < re) {
                double w = weight(nb.dist, re);
                V3 d     = sub(o.vel[nb.id], o.vel[i]);
                t.x += d.x * w;
                t.y += d.y * w;
                t.z += d.z * w;
            }
        }
        t.x *= a; t.y *= a; t.z *= a;
        o.acc[i].x += t.x; o.acc[i].y += t.y; o.acc[i].z += t.z;
    }
}

// MPS::checkBoundaryViolation (src/mps.cpp:643-667)
void checkBoundaryViolation(Oracle& o, int i) {
    const V3& p = o.pos[i];
    bool in = true;
    if (p.x < o.s.domain_min[0] || o.s.domain_max[0] < p.x) in = false;
    if (p.y < o.s.domain_min[1] || o.s.domain_max[1] < p.y) in = false;
    if (p.z < o.s.domain_min[2] || o.s.domain_max[2] < p.z) in = false;
    if (!in) {
        o.debugLogCount++;
        o.type[i] = MPSB_GHOST;
    }
}

// MPS::moveParticles (src/mps.cpp:205-215)
void moveParticles(Oracle& o) {
    const double dt = o.s.dt;
    for (int i = 0; i < o.n; i++) {
        if (o.type[i] == MPSB_FLUID) {
            o.vel[i].x += o.acc[i].x * dt; o.vel[i].y += o.acc[i].y * dt; o.vel[i].z += o.acc[i].z * dt;
            o.pos[i].x += o.vel[i].x * dt; o.pos[i].y += o.vel[i].y * dt; o.pos[i].z += o.vel[i].z * dt;
            checkBoundaryViolation(o, i);
        }
        o.acc[i] = {0, 0, 0};
    }
}

// Particle::inverseDensity (src/particle.cpp:19-29)
inline double inverseDensity(const Oracle& o, int i) { return o.type[i] == MPSB_FLUID ? 1 / o.s.density : 0.0; }

// MPS::collision (src/mps.cpp:217-250): sequential, in place; list distances are stale, normals live
void collision(Oracle& o) {
    for (int i = 0; i < o.n; i++) {
        if (o.type[i] != MPSB_FLUID) continue;
        for (const Nb& nb : o.nbrs[i]) {
            int j = nb.id;
            if (o.type[j] == MPSB_FLUID && j >= i) continue;
            if (nb.dist < o.s.collision_distance) {
                double invMi = inverseDensity(o, i), invMj = inverseDensity(o, j);
                double reducedMass = 1.0 / (invMi + invMj);
                V3 d       = sub(o.pos[j], o.pos[i]);
                double z   = dot(d, d);
                V3 nrm     = d; // Eigen normalized(): divide by sqrt(z) when z > 0 (Dot.h:124-134)
                if (z > 0) {
                    double sz = std::sqrt(z);
                    nrm = {d.x / sz, d.y / sz, d.z / sz};
                }
                double rel = dot(sub(o.vel[j], o.vel[i]), nrm);
                if (rel < 0.0) {
                    double impulse = -(1.0 + o.s.restitution) * rel * reducedMass;
                    double ci = impulse * invMi, cj = impulse * invMj;
                    o.vel[i].x -= ci * nrm.x; o.vel[i].y -= ci * nrm.y; o.vel[i].z -= ci * nrm.z;
                    o.vel[j].x += cj * nrm.x; o.vel[j].y += cj * nrm.y; o.vel[j].z += cj * nrm.z;
                    double depth = o.s.collision_distance - nb.dist;
                    double posImpulse = depth * reducedMass;
                    double pi = posImpulse * invMi, pj = posImpulse * invMj;
                    o.pos[i].x -= pi * nrm.x; o.pos[i].y -= pi * nrm.y; o.pos[i].z -= pi * nrm.z;
                    o.pos[j].x += pj * nrm.x; o.pos[j].y += pj * nrm.y; o.pos[j].z += pj * nrm.z;
                }
            }
        }
    }
}

// MPS::setBoundaryCondition (src/mps.cpp:262-278)
void setBoundaryCondition(Oracle& o) {
    double n0 = o.s.n0_surface, beta = o.s.surface_threshold;
    for (int i = 0; i < o.n; i++) {
        if (o.type[i] == MPSB_GHOST || o.type[i] == MPSB_DUMMY_WALL)
            o.bc[i] = MPSB_BC_IGNORED;
        else if (numberDensity(o, i, o.s.re_surface) < beta * n0)
            o.bc[i] = MPSB_BC_SURFACE;
        else
            o.bc[i] = MPSB_BC_INNER;
    }
}

// MPS::setSourceTerm (src/mps.cpp:280-320)
void setSourceTerm(Oracle& o) {
    const double re = o.s.re_pressure, n0 = o.s.n0_pressure, dt = o.s.dt;
    for (int i = 0; i < o.n; i++) {
        if (o.bc[i] != MPSB_BC_INNER) {
            o.rhs[i] = 0.0;
            continue;
        }
        if (o.s.higher_order_source_on) {
            const double gamma = o.s.higher_order_gamma;
            o.rhs[i] = gamma * (1.0 / (dt * dt)) * (o.ratio[i] - 1.0);
            for (const Nb& nb : o.nbrs[i]) {
                int j = nb.id;
                if (o.bc[j] == MPSB_BC_IGNORED) continue;
                V3 r = sub(o.pos[j], o.pos[i]), u = sub(o.vel[j], o.vel[i]);
                double dist = nb.dist;
                if (dist < re) o.rhs[i] -= (1.0 / (n0 * dt)) * re * dot(r, u) / (dist * dist * dist);
            }
        } else {
            const double gamma = o.s.relaxation_coefficient;
            const double n = numberDensity(o, i, re);
            o.rhs[i] = gamma * (1.0 / (dt * dt)) * (n - n0) / n0;
        }
    }
}

// MPS::setMatrix (src/mps.cpp:322-360) + Eigen setFromTriplets (rows sorted by column)
void setMatrix(Oracle& o) {
    const double re = o.s.re_pressure, n0 = o.s.n0_pressure, lambda = o.s.lambda_pressure;
    const double D = (double) o.s.dim;
    o.rowStart.assign(o.n + 1, 0);
    o.col.clear();
    o.val.clear();
    std::vector<std::pair<int, double>> row;
    for (int i = 0; i < o.n; i++) {
        o.rowStart[i] = (int64_t) o.col.size();
        if (o.bc[i] != MPSB_BC_INNER) continue;
        row.clear();
        double cii = 0.0;
        for (const Nb& nb : o.nbrs[i]) {
            int j = nb.id;
            if (o.bc[j] == MPSB_BC_IGNORED) continue;
            if (nb.dist < re) {
                double a = (2.0 * D / (lambda * n0)) * weight(nb.dist, re) / o.s.density;
                cii += a;
                if (o.bc[j] == MPSB_BC_INNER) row.push_back({j, -a});
            }
        }
        if (o.s.quasi_compressibility_on) cii += o.s.compressibility / (o.s.dt * o.s.dt);
        row.push_back({i, cii});
        std::sort(row.begin(), row.end(), [](auto& a, auto& b) { return a.first < b.first; });
        for (auto& e : row) {
            o.col.push_back(e.first);
            o.val.push_back(e.second);
        }
    }
    o.rowStart[o.n] = (int64_t) o.col.size();
}

// MPS::ensureDirichletBoundaryConnection (src/mps.cpp:362-415)
void ensureDirichlet(Oracle& o) {
    for (int i = 0; i < o.n; i++) o.connected[i] = o.bc[i] == MPSB_BC_SURFACE;
    for (int s = 0; s < o.n; s++) {
        if (o.bc[s] != MPSB_BC_SURFACE) continue;
        std::queue<int> q;
        q.push(s);
        while (!q.empty()) {
            int i = q.front();
            q.pop();
            for (const Nb& nb : o.nbrs[i])
                if (nb.dist < o.s.re_pressure && !o.connected[nb.id]) {
                    o.connected[nb.id] = 1;
                    q.push(nb.id);
                }
        }
    }
    for (int i = 0; i < o.n; i++)
        if (!o.connected[i] && o.bc[i] == MPSB_BC_INNER) {
            o.debugLogCount++;
            for (int64_t k = o.rowStart[i]; k < o.rowStart[i + 1]; k++)
                if (o.col[k] == i) o.val[k] *= 2.0;
        }
}

// Eigen's vectorised sum of a length-n expression e(i): SSE2 packets of 2, two interleaved packet
// accumulators, then lane0+lane1, then the scalar tail (Eigen/src/Core/Redux.h,
// LinearVectorizedTraversal / NoUnrolling; alignedStart = 0 for expression evaluators).
template <class F>
double eigenSum(int n, F e) {
    const int ps = 2;
    int alignedSize2 = (n / (2 * ps)) * (2 * ps), alignedSize = (n / ps) * ps;
    double res;
    if (alignedSize) {
        double a0 = e(0), a1 = e(1);
        if (alignedSize > ps) {
            double b0 = e(2), b1 = e(3);
            for (int i = 2 * ps; i < alignedSize2; i += 2 * ps) {
                a0 += e(i); a1 += e(i + 1);
                b0 += e(i + 2); b1 += e(i + 3);
            }
            a0 = a0 + b0; a1 = a1 + b1;
            if (alignedSize > alignedSize2) {
                a0 += e(alignedSize2); a1 += e(alignedSize2 + 1);
            }
        }
        res = a0 + a1;
        for (int i = alignedSize; i < n; i++) res += e(i);
    } else {
        res = e(0);
        for (int i = 1; i < n; i++) res += e(i);
    }
    return res;
}

// Eigen ConjugateGradient<SparseMatrix<double>, Lower, DiagonalPreconditioner> exactly as the
// reference uses it (src/mps.cpp:452-457): conjugate_gradient() (ConjugateGradient.h:28-91),
// x0 = 0, tol = DBL_EPSILON, maxIters = 2n (IterativeSolverBase.h:281-284,400-413), Jacobi with
// invdiag = 1 where the diagonal is missing or 0 (BasicPreconditioners.h:65-79), and
// A.selfadjointView<Lower>() * p on the ColMajor copy (SparseSelfAdjointView.h:283-340).
void cgSolve(Oracle& o, std::vector<double>& x) {
    const int n = o.n;
    // ColMajor copy = transpose of the row CSR (columns hold ascending row indices)
    std::vector<int64_t> cs(n + 1, 0);
    for (int c : o.col) cs[c + 1]++;
    for (int j = 0; j < n; j++) cs[j + 1] += cs[j];
    std::vector<int> ri(o.col.size());
    std::vector<double> cv(o.col.size());
    {
        std::vector<int64_t> fill(cs.begin(), cs.end() - 1);
        for (int i = 0; i < n; i++)
            for (int64_t k = o.rowStart[i]; k < o.rowStart[i + 1]; k++) {
                int64_t d = fill[o.col[k]]++;
                ri[d] = i;
                cv[d] = o.val[k];
            }
    }
    std::vector<double> invdiag(n, 1.0);
    for (int j = 0; j < n; j++)
        for (int64_t k = cs[j]; k < cs[j + 1]; k++)
            if (ri[k] == j && cv[k] != 0.0) invdiag[j] = 1.0 / cv[k];

    auto spmv = [&](const std::vector<double>& v, std::vector<double>& res) {
        std::fill(res.begin(), res.end(), 0.0);
        for (int j = 0; j < n; j++) {
            int64_t k = cs[j], e = cs[j + 1];
            while (k < e && ri[k] < j) k++;
            if (k < e && ri[k] == j) {
                res[j] += cv[k] * v[j];
                k++;
            }
            double rhs_j = v[j], res_j = 0.0;
            for (; k < e; k++) {
                res_j += cv[k] * v[ri[k]];
                res[ri[k]] += cv[k] * rhs_j;
            }
            res[j] += res_j;
        }
    };

    x.assign(n, 0.0);
    const std::vector<double>& b = o.rhs;
    std::vector<double> r(b), p(n), z(n), tmp(n);
    const double tol = DBL_EPSILON;
    int maxIters = 2 * n;
    double rhsNorm2 = eigenSum(n, [&](int i) { return b[i] * b[i]; });
    o.cgIterations = 0;
    o.cgError      = 0;
    if (n == 0 || rhsNorm2 == 0) return;
    double threshold = std::max(tol * tol * rhsNorm2, DBL_MIN);
    double residualNorm2 = eigenSum(n, [&](int i) { return r[i] * r[i]; });
    if (residualNorm2 < threshold) {
        o.cgError = std::sqrt(residualNorm2 / rhsNorm2);
        return;
    }
    for (int i = 0; i < n; i++) p[i] = invdiag[i] * r[i];
    double absNew = eigenSum(n, [&](int i) { return r[i] * p[i]; });
    int it = 0;
    while (it < maxIters) {
        spmv(p, tmp);
        double alpha = absNew / eigenSum(n, [&](int i) { return p[i] * tmp[i]; });
        for (int i = 0; i < n; i++) x[i] += alpha * p[i];
        for (int i = 0; i < n; i++) r[i] -= alpha * tmp[i];
        residualNorm2 = eigenSum(n, [&](int i) { return r[i] * r[i]; });
        if (residualNorm2 < threshold) break;
        for (int i = 0; i < n; i++) z[i] = invdiag[i] * r[i];
        double absOld = absNew;
        absNew        = eigenSum(n, [&](int i) { return r[i] * z[i]; });
        double beta   = absNew / absOld;
        for (int i = 0; i < n; i++) p[i] = z[i] + beta * p[i];
        it++;
    }
    o.cgError      = std::sqrt(residualNorm2 / rhsNorm2);
    o.cgIterations = it;
}

// MPS::solvePoissonEquation (src/mps.cpp:417-470); the second, redundant solve is not repeated
void solvePoisson(Oracle& o) {
    std::vector<double> x;
    cgSolve(o, x);
    for (int i = 0; i < o.n; i++) o.pressure[i] = x[i];
}

// MPS::removeNegativePressure (src/mps.cpp:472-479)
void removeNegativePressure(Oracle& o) {
    for (int i = 0; i < o.n; i++)
        if (o.pressure[i] < 0.0) o.pressure[i] = 0.0;
}

// MPS::setMinimumPressure (src/mps.cpp:481-503)
void setMinimumPressure(Oracle& o) {
    for (int i = 0; i < o.n; i++) {
        if (o.bc[i] == MPSB_BC_IGNORED) continue;
        o.minp[i] = o.pressure[i];
        for (const Nb& nb : o.nbrs[i]) {
            if (o.type[nb.id] == MPSB_DUMMY_WALL) continue;
            if (nb.dist < o.s.re_pressure)
                if (o.minp[i] > o.pressure[nb.id]) o.minp[i] = o.pressure[nb.id];
        }
    }
}

// MPS::calcPressureGradient (src/mps.cpp:505-533)
void calcPressureGradient(Oracle& o) {
    const double re = o.s.re_pressure, n0 = o.s.n0_pressure;
    for (int i = 0; i < o.n; i++) {
        if (o.type[i] != MPSB_FLUID) continue;
        V3 g{0, 0, 0};
        for (const Nb& nb : o.nbrs[i]) {
            int j = nb.id;
            if (o.type[j] == MPSB_DUMMY_WALL) continue;
            double dist = nb.dist;
            if (dist < re) {
                V3 d      = sub(o.pos[j], o.pos[i]);
                double dp = o.pressure[j] - o.minp[i];
                double w  = weight(dist, re);
                double d2 = dist * dist;
                g.x += d.x * dp * w / d2;
                g.y += d.y * dp * w / d2;
                g.z += d.z * dp * w / d2;
            }
        }
        double c = (double) o.s.dim / n0;
        g.x *= c; g.y *= c; g.z *= c;
        o.acc[i] = {-1.0 * g.x / o.s.density, -1.0 * g.y / o.s.density, -1.0 * g.z / o.s.density};
    }
}

// MPS::moveParticlesWithPressureGradient (src/mps.cpp:535-545); acceleration is NOT cleared (B-3)
void moveParticlesWithPressureGradient(Oracle& o) {
    const double dt = o.s.dt;
    for (int i = 0; i < o.n; i++) {
        if (o.type[i] != MPSB_FLUID) continue;
        o.vel[i].x += o.acc[i].x * dt; o.vel[i].y += o.acc[i].y * dt; o.vel[i].z += o.acc[i].z * dt;
        o.pos[i].x += o.acc[i].x * dt * dt; o.pos[i].y += o.acc[i].y * dt * dt; o.pos[i].z += o.acc[i].z * dt * dt;
        checkBoundaryViolation(o, i);
    }
}

// MPS::getCourantNumber (src/mps.cpp:547-571)
double courant(Oracle& o) {
    double m = 0.0;
    for (int i = 0; i < o.n; i++) {
        if (o.type[i] != MPSB_FLUID) continue;
        m = std::max(m, (norm(o.vel[i]) * o.s.dt) / o.s.particle_distance);
    }
    if (m > o.s.cfl_condition) o.debugLogCount++;
    return m;
}

// MPS::calcPressure (src/mps.cpp:252-260)
void calcPressure(Oracle& o) {
    setBoundaryCondition(o);
    setSourceTerm(o);
    setMatrix(o);
    ensureDirichlet(o);
    solvePoisson(o);
    removeNegativePressure(o);
    setMinimumPressure(o);
}

// MPS::stepForward (src/mps.cpp:111-130)
double stepForward(Oracle& o, bool isTimeToExport) {
    setNeighbors(o);
    calcGravity(o);
    calcViscosity(o);
    moveParticles(o);
    setNeighbors(o);
    collision(o);
    setNeighbors(o);
    calcPressure(o);
    calcPressureGradient(o);
    moveParticlesWithPressureGradient(o);
    if (isTimeToExport || o.s.higher_order_source_on) setNumberDensity(o);
    return courant(o);
}

} // namespace

extern "C" {

// MPS::loadInitialState after the CSV import (src/mps.cpp:99-106)
void* orc_create(const mpsb_settings* s, int64_t n, const int32_t* type, const double* pos, const double* vel) {
    auto* o = new Oracle;
    o->s    = *s;
    o->n    = (int) n;
    o->type.assign(type, type + n);
    o->bc.assign(n, MPSB_BC_IGNORED);
    o->connected.assign(n, 0);
    o->pos.resize(n); o->vel.resize(n); o->acc.assign(n, {0, 0, 0});
    for (int64_t i = 0; i < n; i++) {
        o->pos[i] = {pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]};
        o->vel[i] = {vel[3 * i], vel[3 * i + 1], vel[3 * i + 2]};
    }
    o->pressure.assign(n, 0.0); o->minp.assign(n, 0.0); o->ratio.assign(n, 0.0);
    o->nbrs.resize(n);
    o->rhs.assign(n, 0.0);
    o->rowStart.assign(n + 1, 0);
    bucketInit(*o);
    setNumberDensity(*o);
    return o;
}
void orc_destroy(void* h) { delete static_cast<Oracle*>(h); }
int64_t orc_num_particles(void* h) { return static_cast<Oracle*>(h)->n; }
int orc_debug_log_count(void* h) { return static_cast<Oracle*>(h)->debugLogCount; }
void orc_get_settings(void* h, mpsb_settings* s) { *s = static_cast<Oracle*>(h)->s; }
void orc_set_pressure_options(void* h, int qc, double alpha, int ho, double gamma, double relax) {
    mpsb_settings& s = static_cast<Oracle*>(h)->s;
    s.quasi_compressibility_on = qc; s.compressibility = alpha;
    s.higher_order_source_on = ho; s.higher_order_gamma = gamma;
    s.relaxation_coefficient = relax;
}

void orc_get_state(void* h, const mpsb_fields* f) {
    Oracle& o = *static_cast<Oracle*>(h);
    for (int i = 0; i < o.n; i++) {
        if (f->type) f->type[i] = o.type[i];
        if (f->position) { f->position[3 * i] = o.pos[i].x; f->position[3 * i + 1] = o.pos[i].y; f->position[3 * i + 2] = o.pos[i].z; }
        if (f->velocity) { f->velocity[3 * i] = o.vel[i].x; f->velocity[3 * i + 1] = o.vel[i].y; f->velocity[3 * i + 2] = o.vel[i].z; }
        if (f->acceleration) { f->acceleration[3 * i] = o.acc[i].x; f->acceleration[3 * i + 1] = o.acc[i].y; f->acceleration[3 * i + 2] = o.acc[i].z; }
        if (f->pressure) f->pressure[i] = o.pressure[i];
        if (f->minimum_pressure) f->minimum_pressure[i] = o.minp[i];
        if (f->number_density_ratio) f->number_density_ratio[i] = o.ratio[i];
        if (f->boundary_condition) f->boundary_condition[i] = o.bc[i];
    }
}
void orc_set_state(void* h, const mpsb_fields* f) {
    Oracle& o = *static_cast<Oracle*>(h);
    for (int i = 0; i < o.n; i++) {
        if (f->type) o.type[i] = f->type[i];
        if (f->position) o.pos[i] = {f->position[3 * i], f->position[3 * i + 1], f->position[3 * i + 2]};
        if (f->velocity) o.vel[i] = {f->velocity[3 * i], f->velocity[3 * i + 1], f->velocity[3 * i + 2]};
        if (f->acceleration) o.acc[i] = {f->acceleration[3 * i], f->acceleration[3 * i + 1], f->acceleration[3 * i + 2]};
        if (f->pressure) o.pressure[i] = f->pressure[i];
        if (f->minimum_pressure) o.minp[i] = f->minimum_pressure[i];
        if (f->number_density_ratio) o.ratio[i] = f->number_density_ratio[i];
        if (f->boundary_condition) o.bc[i] = f->boundary_condition[i];
    }
}

double orc_phase(void* h, int phase) {
    Oracle& o = *static_cast<Oracle*>(h);
    switch (phase) {
    case 0: setNeighbors(o); break;
    case 1: calcGravity(o); break;
    case 2: calcViscosity(o); break;
    case 3: moveParticles(o); break;
    case 4: collision(o); break;
    case 5: setBoundaryCondition(o); break;
    case 6: setSourceTerm(o); break;
    case 7: setMatrix(o); break;
    case 8: ensureDirichlet(o); break;
    case 9: solvePoisson(o); break;
    case 10: removeNegativePressure(o); break;
    case 11: setMinimumPressure(o); break;
    case 12: calcPressureGradient(o); break;
    case 13: moveParticlesWithPressureGradient(o); break;
    case 14: setNumberDensity(o); break;
    case 15: return courant(o);
    default: return -1.0;
    }
    return 0.0;
}
double orc_step(void* h, int exp) { return stepForward(*static_cast<Oracle*>(h), exp != 0); }

int64_t orc_get_neighbors(void* h, int64_t* row_start, int32_t* ids, double* dist, int64_t capacity) {
    Oracle& o = *static_cast<Oracle*>(h);
    int64_t k = 0;
    for (int i = 0; i < o.n; i++) {
        if (row_start) row_start[i] = k;
        if (o.type[i] == MPSB_GHOST) continue;
        for (const Nb& nb : o.nbrs[i]) {
            if (ids && k < capacity) { ids[k] = nb.id; dist[k] = nb.dist; }
            k++;
        }
    }
    if (row_start) row_start[o.n] = k;
    return k;
}
int64_t orc_get_ppe(void* h, int64_t* row_start, int32_t* col, double* val, double* rhs, int64_t capacity) {
    Oracle& o = *static_cast<Oracle*>(h);
    int64_t nnz = (int64_t) o.col.size();
    if (row_start) std::copy(o.rowStart.begin(), o.rowStart.end(), row_start);
    if (col)
        for (int64_t k = 0; k < nnz && k < capacity; k++) { col[k] = o.col[k]; val[k] = o.val[k]; }
    if (rhs) std::copy(o.rhs.begin(), o.rhs.end(), rhs);
    return nnz;
}
int orc_cg_solve(void* h, double* x, double* error, int* info) {
    Oracle& o = *static_cast<Oracle*>(h);
    std::vector<double> v;
    cgSolve(o, v);
    if (x) std::copy(v.begin(), v.end(), x);
    if (error) *error = o.cgError;
    if (info) *info = o.cgError <= DBL_EPSILON ? 0 : 2; // Eigen: Success / NoConvergence
    return o.cgIterations;
}
void orc_get_dirichlet_connected(void* h, int32_t* out) {
    Oracle& o = *static_cast<Oracle*>(h);
    std::copy(o.connected.begin(), o.connected.end(), out);
}

} // extern "C"
