#include "simulation.hpp"

#include "csv_io.hpp"
#include "exporter.hpp"
#include "log.hpp"
#include "rendezvous.hpp"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <format>
#include <iostream>
#include <string>

namespace fs = std::filesystem;
using std::chrono::duration_cast;
using std::chrono::milliseconds;
using std::chrono::seconds;
using std::chrono::system_clock;

namespace {
// "{:%T}" of a seconds count: HH:MM:SS (hours not wrapped)
std::string hms(long long s) {
    if (s < 0) s = 0;
    char buf[48];
    std::snprintf(buf, sizeof buf, "%02lld:%02lld:%02lld", s / 3600, (s / 60) % 60, s % 60);
    return buf;
}
// "{:%S}" of a milliseconds count: seconds of the minute with three decimals
std::string secondsOfMinute(long long ms) {
    char buf[32];
    std::snprintf(buf, sizeof buf, "%02lld.%03lld", (ms / 1000) % 60, ms % 1000);
    return buf;
}
} // namespace

void Simulation::run(const fs::path& inputYamlPath) {
    // multi-GPU (one process per GPU): every rank steps the same simulation; only rank 0 writes files and logs
    writer = rendezvous::fromEnvironment().rank == 0;
    if (writer) {
        createResultDirectory(inputYamlPath);
        prepareLogFile();
    } else {
        mpslog::Logger::get().setConsole(false);
        quiet = true;
    }

    mps  = MPS(inputYamlPath);
    time = stateIn.empty() ? mps.loadInitialState() : mps.loadInitialStateFromFile(stateIn);
    if (writer) exportParticles(mps.particles);

    mpslog::info("START SIMULATION");
    simulationStartTime = system_clock::now();
    // `time` is ACCUMULATED in double, not computed as step*dt: the step count and the export cadence
    // of the reference depend on it (SURVEY B-14: hydrostatic runs 999 steps, not 1000).
    while (time + mps.settings.dt <= mps.settings.finishTime) {
        if (maxSteps > 0 && timeStep >= maxSteps) break;
        auto timeStepStartTime   = system_clock::now();
        int initialDebugLogCount = mps.debugLogCount;

        time += mps.settings.dt;
        const bool exportNow = isTimeToExport();
        double courantNumber = mps.stepForward(exportNow);
        if (exportNow && writer) exportParticles(mps.particles);

        auto timeStepEndTime = system_clock::now();
        if (writer) timeStepReport(timeStepStartTime, timeStepEndTime, courantNumber, mps.debugLogCount > initialDebugLogCount);
        timeStep++;
    }
    simulationEndTime = system_clock::now();
    if (!stateOut.empty()) { // the download is collective with several ranks (every rank joins); rank 0 alone writes
        mps.downloadParticles();
        if (writer && !mps.saveState(stateOut, time, false)) mpslog::error("Could not write the state file: " + stateOut.string());
    }
    endSimulation();
}

void Simulation::createResultDirectory(const fs::path& inputYamlPath) {
    // <yaml dir>/result is REPLACED, as the reference does (src/simulation.cpp:66-77, SURVEY B-15)
    resultDirectory = inputYamlPath.parent_path() / "result";
    if (fs::exists(resultDirectory)) fs::remove_all(resultDirectory);
    fs::create_directory(resultDirectory);
    fs::create_directory(resultDirectory / "csv");
    fs::create_directory(resultDirectory / "vtu");
    fs::copy(inputYamlPath, resultDirectory / inputYamlPath.filename());
}

void Simulation::prepareLogFile() {
    mpslog::Logger::get().setConsole(true);
    mpslog::Logger::get().openFile((resultDirectory / "execution.log").string());

    fs::path path = resultDirectory / "time_step_report.csv";
    timeStepReportFile.open(path);
    if (!timeStepReportFile.is_open()) mpslog::error("Could not open the log file: " + path.string());
    timeStepReportFile << "Time Step,Dt (s),Time Before (s),Time After (s),Finish Time (s),Elapsed Time (h:m:s),Elapsed Time (s),"
                          "Remaining Time (h:m:s),Average Time per Time Step (s),Time for This Time Step (s),Number of Output Files,"
                          "Courant Number";
    if (extendedReport) timeStepReportFile << ",CG Iterations,PPE Solve (ms),GPU Step (ms),PPE Unknowns";
    timeStepReportFile << "\n";
}

void Simulation::endSimulation() {
    auto total = duration_cast<seconds>(simulationEndTime - simulationStartTime);
    mpslog::info("END SIMULATION");
    if (mps.debugLogCount > 0)
        mpslog::info("There were " + std::to_string(mps.debugLogCount) + " debug-level logs. Check 'execution.log'.");
    else
        mpslog::info("There were no debug-level logs.");
    mpslog::info("Total Simulation Time = " + hms(total.count()));
    timeStepReportFile.close();
    mpslog::Logger::get().closeFile();
}

void Simulation::timeStepReport(
    const system_clock::time_point& timeStepStartTime, const system_clock::time_point& timeStepEndTime, const double& courantNumber,
    const bool& isDebugLogAdded
) {
    auto elapsed = duration_cast<seconds>(timeStepEndTime - simulationStartTime);
    double average = 0;
    if (timeStep != 0) average = duration_cast<milliseconds>(timeStepEndTime - simulationStartTime).count() / (double) (1000 * timeStep);
    // the reference's estimate, kept as is (SURVEY B-13)
    long long remain = (long long) (int) (((mps.settings.finishTime - time) / time) * average * timeStep);
    auto last        = duration_cast<milliseconds>(timeStepEndTime - timeStepStartTime);

    const std::string timeBefore = std::format("{:.3f}", time - mps.settings.dt), timeAfter = std::format("{:.3f}", time);
    const std::string elapsedText = hms(elapsed.count()), remainText = hms(remain), averageText = std::format("{:.3f}", average);
    const std::string lastText = secondsOfMinute(last.count()), courantText = std::format("{:.2f}", courantNumber);

    if (!quiet)
        std::cout << std::format(
                         "{}: dt={}s  t={}s->{}s  fin={}s  elapsed={}  remain={}  ave={}s/step  last={}s/step  out={}files  Courant={}",
                         timeStep, mps.settings.dt, timeBefore, timeAfter, mps.settings.finishTime, elapsedText, remainText, averageText, lastText,
                         outFileNum, courantText
                     )
                  << std::endl;

    Exporter spelling; // float spelling of the report follows the particle CSVs
    timeStepReportFile << timeStep << ',' << csv_io::formatFloat(mps.settings.dt, spelling.floatStyle) << ',' << timeBefore << ',' << timeAfter << ','
                       << csv_io::formatFloat(mps.settings.finishTime, spelling.floatStyle) << ',' << elapsedText << ',' << elapsed.count() << ','
                       << remainText << ',' << averageText << ',' << lastText << ',' << outFileNum << ',' << courantText;
    if (extendedReport)
        timeStepReportFile << ',' << mps.lastStep.cgIterations << ',' << std::format("{:.3f}", mps.lastStep.msSolve) << ','
                           << std::format("{:.3f}", mps.lastStep.msTotal) << ',' << mps.lastStep.nInner;
    timeStepReportFile << '\n';

    if (isDebugLogAdded) mpslog::debug(std::format("Time Step:{} Time:{}s->{}s", timeStep, timeBefore, timeAfter));
}

bool Simulation::isTimeToExport() {
    double outputInterval = mps.settings.outputInterval;
    double remainder      = std::fmod(time, outputInterval);
    remainder             = std::min(remainder, outputInterval - remainder);
    return remainder < mps.settings.dt / 10.0;
}

void Simulation::exportParticles(const std::vector<Particle>& particles) {
    Exporter exporter;
    exporter.toCsv(resultDirectory / std::format("csv/output_{:04}.csv", outFileNum), particles, time);
    exporter.toVtu(resultDirectory / std::format("vtu/output_{:04}.vtu", outFileNum), particles, time);
    outFileNum++;
}
