// csv_float_probe.cpp — TEST INFRASTRUCTURE. Prints each argument (a double) the way the reference's CSV writer
// spells it in a gcc build: csv::internals::to_string<double> of the csv-parser header vendored by the reference
// (submodules/csv-parser/single_include/csv.hpp:6636-6681), which this file only includes. Built into
// oracle/_ref/csv_float.exe by oracle/Makefile; tests/test_host.py compares mps_b200/host/csv_io.hpp with it.
#include "csv.hpp"

#include <cstdio>
#include <cstdlib>

int main(int argc, char** argv) {
    for (int i = 1; i < argc; i++) std::printf("%s\n", csv::internals::to_string(std::strtod(argv[i], nullptr)).c_str());
    return 0;
}
