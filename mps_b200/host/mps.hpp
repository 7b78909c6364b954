// mps.hpp — host mirror of the reference's `class MPS` (reference src/mps.hpp:10-71) over the C ABI of
// libmps_b200.so (include/mps_b200.h). Public members keep the reference's names and meaning:
//   settings, particles, debugLogCount, MPS(), MPS(yaml), loadInitialState(), stepForward(bool).
// What the reference holds privately (Bucket, Eigen matrix, the phase functions of src/mps.cpp) is
// replaced by ONE opaque device handle; every phase runs on the GPU. There is no CPU path: without a
// CUDA device loadInitialState() logs the library's error and exits (SURVEY §8b error conventions).
#pragma once

#include "particle.hpp"
#include "settings.hpp"

#include <cstdint>
#include <filesystem>
#include <memory>
#include <vector>

struct mpsb_handle;
struct mpsb_step_stats;

class RefValues {
public:
    double initialNumberDensity = 0;
    double lambda               = 0;

    RefValues() = default;
    // n0 and lambda over the +-4 lattice with the reference's loop order and operation order
    // (reference src/mps.cpp:132-163) so both are bit-identical to the reference's.
    RefValues(const int& dim, const double& particleDistance, const double& effectiveRadius);
};

class MPS {
public:
    Settings settings;
    std::vector<Particle> particles;
    int debugLogCount = 0;

    MPS() = default;
    MPS(const std::filesystem::path& inputYamlPath);

    double loadInitialState();                     // reads the input CSV, creates the device state; returns initial time
    double stepForward(const bool isTimeToExport); // one MPS step on the GPU; returns the Courant number

    // ---- additions (not in the reference) -------------------------------------------------------
    // `particles` is refreshed from HBM at load and on steps called with isTimeToExport == true (the
    // only times the reference's own driver reads it, src/simulation.cpp:35,48). Callers that read
    // `particles` after other steps call downloadParticles() first, or set syncEveryStep.
    bool syncEveryStep = false;
    void downloadParticles();
    void uploadParticles();   // push host edits of type/position/velocity back to the device
    void downloadNeighbors(); // fill Particle::neighbors from the device lists (debugging aid; O(N k) host memory)
    struct StepReport {       // mpsb_step_stats of the last step (include/mps_b200.h)
        int cgIterations = 0, nInner = 0, nSurface = 0, nnz = 0, nActive = 0, nCollisions = 0;
        double cgRelResidual = 0, msTotal = 0, msSolve = 0, msNeighbors = 0;
    } lastStep;
    // Initial state from arrays instead of settings.inputCsvPath (bench / generated cases: CSV cannot
    // carry fp64 positions). Same effect as loadInitialState(); returns `time`.
    double loadInitialStateFromArrays(
        double time, std::int64_t n, const std::int32_t* type, const double* position3, const double* velocity3
    );
    double loadParticlesOnly() { return importInitialCondition(); } // CSV -> particles, no device (host tests, converters)
    // Binary state files (state_io.hpp): full-precision checkpoints. Loading one restores type, position,
    // velocity and the derived fields a following step can depend on (pressure, number-density ratio, flags),
    // so save + load + continue equals an uninterrupted run bit for bit. Returns the file's time.
    double loadInitialStateFromFile(const std::filesystem::path& statePath);
    // downloads first unless told the particles are current (multi-GPU: the download is collective, the write is rank 0's)
    bool saveState(const std::filesystem::path& statePath, double time, bool download = true);
    int device = 0; // CUDA device ordinal used by loadInitialState (MPSB_DEVICE overrides)
    // Multi-GPU: one process per GPU. With WORLD_SIZE > 1 in the environment (rendezvous.hpp) the ranks split the
    // PPE solve; every rank holds the whole particle set and returns identical results.
    int rank = 0, worldSize = 1;

    const RefValues& refValuesPressure() const { return refValues.pressure; }
    const RefValues& refValuesViscosity() const { return refValues.viscosity; }
    const RefValues& refValuesSurfaceDetection() const { return refValues.surfaceDetection; }

private:
    struct {
        RefValues pressure;
        RefValues viscosity;
        RefValues surfaceDetection;
    } refValues;

    // shared_ptr: the reference's driver copy-assigns `mps = MPS(path)` (src/simulation.cpp:33); copies
    // share the device state, the last owner destroys it.
    std::shared_ptr<mpsb_handle> handle;
    // page-locked staging for downloadParticles (mpsb_host_alloc): allocated once, DMA target of mpsb_download
    std::shared_ptr<void> staging;
    std::size_t stagingBytes = 0;

    double importInitialCondition(); // reference src/mps.cpp:38-94
    void createDeviceState(bool restoreDerived = false);
    void reportEvents(const mpsb_step_stats& stats);
    [[noreturn]] void fail(const char* what, int status);
};
