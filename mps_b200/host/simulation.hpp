// simulation.hpp — host mirror of the reference's driver (reference src/simulation.hpp:11-39,
// src/simulation.cpp:29-245): result directory layout, execution.log, time_step_report.csv (the
// reference's 12 columns first), console line per step, export cadence.
#pragma once

#include "mps.hpp"
#include "particle.hpp"

#include <chrono>
#include <filesystem>
#include <fstream>
#include <vector>

class Simulation {
public:
    void run(const std::filesystem::path& inputYamlPath);

    // additions: extra report columns (CG iterations, PPE solve ms, GPU step ms) after the reference's 12;
    // quiet console; step cap for smoke runs (0 = none). Set from the CLI / environment by main().
    bool extendedReport = true;
    bool quiet          = false;
    int maxSteps        = 0;
    std::filesystem::path stateIn, stateOut; // binary checkpoint to start from / to write at the end (state_io.hpp)

private:
    MPS mps;
    bool writer = true; // rank 0 of a multi-GPU run (or the only process)

    int timeStep = 0;
    double time  = 0.0;
    std::chrono::system_clock::time_point simulationStartTime, simulationEndTime;

    int outFileNum = 0;
    std::ofstream timeStepReportFile;
    std::filesystem::path resultDirectory;

    void createResultDirectory(const std::filesystem::path& inputYamlPath);
    void prepareLogFile();
    void timeStepReport(
        const std::chrono::system_clock::time_point& timeStepStartTime,
        const std::chrono::system_clock::time_point& timeStepEndTime, const double& courantNumber, const bool& isDebugLogAdded
    );
    bool isTimeToExport();
    void exportParticles(const std::vector<Particle>& particles);
    void endSimulation();
};
