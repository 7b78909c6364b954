// exporter.hpp — host mirror of the reference's Exporter (reference src/exporter.hpp:8-29):
// toCsv / toVtu with the same file layout. SURVEY §8 f-1: each file is formatted into one memory buffer
// and written with a single write() instead of one stream insertion + flush per value.
#pragma once

#include "csv_io.hpp"
#include "particle.hpp"

#include <filesystem>
#include <vector>

class Exporter {
public:
    // Float spelling of the CSV: Fixed6 ("%f", what a clang build of the reference writes and what its
    // shipped input.csv files contain) by default; MPSB_CSV_FLOAT=gcc5 selects the 5-decimal truncated
    // spelling of a gcc build for byte-diffing against such a binary.
    csv_io::FloatStyle floatStyle;
    Exporter();

    void toCsv(const std::filesystem::path& outputFilePath, const std::vector<Particle>& particles, const double& time);
    void toVtu(const std::filesystem::path& outputFilePath, const std::vector<Particle>& particles, const double& time);
};
