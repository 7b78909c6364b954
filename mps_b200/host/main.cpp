// main.cpp — the reference's CLI (reference src/main.cpp:12-42): `mps -s|--setting <settings.yml>`.
// Own argument handling (the reference uses argparse). Extra, optional: --max-steps N, --quiet,
// --plain-report (only the reference's 12 report columns), --state-in / --state-out FILE (binary checkpoints).
#include "simulation.hpp"

#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <iostream>
#include <string>

int main(int argc, char** argv) {
    std::string setting;
    Simulation simulation;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        if ((a == "-s" || a == "--setting") && i + 1 < argc) {
            setting = argv[++i];
        } else if (a.rfind("--setting=", 0) == 0) {
            setting = a.substr(10);
        } else if (a == "--max-steps" && i + 1 < argc) {
            simulation.maxSteps = std::atoi(argv[++i]);
        } else if (a == "--state-in" && i + 1 < argc) {
            simulation.stateIn = argv[++i];
        } else if (a == "--state-out" && i + 1 < argc) {
            simulation.stateOut = argv[++i];
        } else if (a == "--quiet") {
            simulation.quiet = true;
        } else if (a == "--plain-report") {
            simulation.extendedReport = false;
        } else if (a == "-h" || a == "--help") {
            std::cout << "Usage: mps [-h] --setting VAR [--max-steps N] [--state-in FILE] [--state-out FILE] [--quiet] [--plain-report]\n\n"
                         "  -s, --setting  path to YAML file for settings. [required]\n";
            return 0;
        } else {
            std::cout << "Unknown argument: " << a << std::endl;
            return EXIT_FAILURE;
        }
    }
    if (setting.empty()) {
        std::cout << "-s/--setting: required." << std::endl;
        return EXIT_FAILURE;
    }
    if (!std::filesystem::exists(setting)) {
        std::cout << "ERROR: Input YAML file cannot be found in specified path: " + setting << std::endl;
        return EXIT_FAILURE;
    }
    std::cout << "Setting file: " << setting << std::endl;
    simulation.run(std::filesystem::path(setting));
    return 0;
}
