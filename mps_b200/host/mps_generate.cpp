// mps_generate.cpp — case generator: the tank lattice of the reference's inputs/*/main.cpp for any
// particle distance, written as a full-precision binary state (state_io.hpp) and, on request, as the reference's
// input.csv. Named geometries are the reference's three; --nx scales a geometry to nx fluid columns.
//   mps_generate --case dambreak|dambreak_fine|hydrostatic [--nx N | --l0 L] --state out.bin [--csv input.csv]
#include "exporter.hpp"
#include "generator.hpp"
#include "state_io.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

int main(int argc, char** argv) {
    std::string name = "dambreak_fine", state, csv;
    double l0 = 0.0;
    long nx   = 0;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        if (a == "--case" && i + 1 < argc) name = argv[++i];
        else if (a == "--nx" && i + 1 < argc) nx = std::atol(argv[++i]);
        else if (a == "--l0" && i + 1 < argc) l0 = std::strtod(argv[++i], nullptr);
        else if (a == "--state" && i + 1 < argc) state = argv[++i];
        else if (a == "--csv" && i + 1 < argc) csv = argv[++i];
        else {
            std::fprintf(stderr, "usage: mps_generate --case NAME [--nx N | --l0 L] --state FILE [--csv FILE]\n");
            return 2;
        }
    }
    generator::Tank t{};
    // constants of reference inputs/<case>/main.cpp:12-19
    if (name == "dambreak") t = {0.01, {0.0, 0.3}, {0.0, 0.25}, {0.0, 0.1}, {0.0, 0.2}, 1, 2};
    else if (name == "dambreak_fine") t = {0.004, {0.0, 1.0}, {0.0, 0.25}, {0.0, 0.4}, {0.0, 0.2}, 1, 2};
    else if (name == "hydrostatic") t = {0.005, {0.0, 0.2}, {0.0, 0.35}, {0.0, 0.2}, {0.0, 0.3}, 1, 2};
    else {
        std::fprintf(stderr, "unknown case %s\n", name.c_str());
        return 2;
    }
    if (nx > 0) t.l0 = (t.fluidX[1] - t.fluidX[0]) / (double) nx;
    if (l0 > 0.0) t.l0 = l0;
    try {
        std::vector<Particle> ps = generator::lattice(t, 1000.0);
        if (!state.empty() && !state_io::write(state, 0.0, ps)) {
            std::fprintf(stderr, "cannot write %s\n", state.c_str());
            return 1;
        }
        if (!csv.empty()) {
            Exporter e;
            e.toCsv(csv, ps, 0.0);
        }
        std::printf("particles %zu l0 %.17g\n", ps.size(), t.l0);
    } catch (const std::exception& ex) {
        std::fprintf(stderr, "ERROR: %s\n", ex.what());
        return 1;
    }
    return 0;
}
