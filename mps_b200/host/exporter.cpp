#include "exporter.hpp"

#include "log.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

namespace {
// operator<<(ostream, double) with default flags == "%g" with precision 6 (reference src/exporter.cpp:83-131)
inline void appendG(std::string& s, double v) {
    char buf[40];
    int n = std::snprintf(buf, sizeof buf, "%g", v);
    s.append(buf, (std::size_t) n);
}
inline void appendInt(std::string& s, long long v) {
    char buf[24];
    int n = std::snprintf(buf, sizeof buf, "%lld", v);
    s.append(buf, (std::size_t) n);
}
bool writeAll(const std::filesystem::path& path, const std::string& data) {
    std::FILE* f = std::fopen(path.string().c_str(), "wb");
    if (!f) return false;
    bool ok = std::fwrite(data.data(), 1, data.size(), f) == data.size();
    return std::fclose(f) == 0 && ok;
}
} // namespace

Exporter::Exporter() : floatStyle(csv_io::FloatStyle::Fixed6) {
    const char* e = std::getenv("MPSB_CSV_FLOAT");
    if (e && std::strcmp(e, "gcc5") == 0) floatStyle = csv_io::FloatStyle::Trunc5Gcc;
}

void Exporter::toCsv(const std::filesystem::path& outFilePath, const std::vector<Particle>& particles, const double& time) {
    std::string out;
    out.reserve(128 + particles.size() * 96);
    out += "Time (s),";
    out += csv_io::formatFloat(time, floatStyle);
    out += "\nNumber of Particles,";
    appendInt(out, (long long) particles.size()); // counts ghosts too, as the reference (SURVEY B-7)
    out += "\nID,Type,Position.x (m),Position.y (m),Position.z (m),Velocity.x (m/s),Velocity.y (m/s),Velocity.z (m/s),"
           "Pressure (Pa),Number Density Ratio\n";
    for (const Particle& pi : particles) {
        if (pi.type == ParticleType::Ghost) continue;
        appendInt(out, pi.id);
        out += ',';
        appendInt(out, static_cast<int>(pi.type));
        const double v[8] = {pi.position.x(), pi.position.y(), pi.position.z(), pi.velocity.x(), pi.velocity.y(), pi.velocity.z(),
                             pi.pressure, pi.numberDensityRatio};
        for (double x : v) {
            out += ',';
            out += csv_io::formatFloat(x, floatStyle);
        }
        out += '\n';
    }
    if (!writeAll(outFilePath, out)) mpslog::error("Could not open target csv file: " + outFilePath.string());
}

void Exporter::toVtu(const std::filesystem::path& outFilePath, const std::vector<Particle>& particles, const double& time) {
    const std::size_t n = particles.size();
    std::string out;
    out.reserve(1024 + n * 160);
    auto arrayBegin = [&](const char* components, const char* type, const char* name) {
        out += "<DataArray NumberOfComponents='";
        out += components;
        out += "' type='";
        out += type;
        out += "' Name='";
        out += name;
        out += "' format='ascii'>\n";
    };
    out += "<?xml version='1.0' encoding='UTF-8'?>\n";
    out += "<VTKFile byte_order='LittleEndian' version='0.1' type = 'UnstructuredGrid'>\n";
    out += "<UnstructuredGrid>\n";
    out += "<FieldData>\n";
    out += "<DataArray type='Float64' Name='Time' NumberOfComponents='1' NumberOfTuples='1' format='ascii'>\n";
    appendG(out, time);
    out += "\n</DataArray>\n</FieldData>\n";
    out += "<Piece NumberOfCells='";
    appendInt(out, (long long) n);
    out += "' NumberOfPoints='";
    appendInt(out, (long long) n);
    out += "'>\n<Points>\n";
    out += "<DataArray NumberOfComponents='3' type='Float64' Name='position' format='ascii'>\n";
    for (const Particle& pi : particles) { // ghosts stay in the VTU (SURVEY B-7)
        appendG(out, pi.position.x());
        out += ' ';
        appendG(out, pi.position.y());
        out += ' ';
        appendG(out, pi.position.z());
        out += '\n';
    }
    out += "</DataArray>\n</Points>\n<PointData>\n";
    arrayBegin("1", "Int32", "Type");
    for (const Particle& pi : particles) {
        appendInt(out, static_cast<int>(pi.type));
        out += '\n';
    }
    out += "</DataArray>\n";
    arrayBegin("3", "Float64", "Velocity");
    for (const Particle& pi : particles) { // z component written as a literal zero, as the reference
        appendG(out, pi.velocity.x());
        out += ' ';
        appendG(out, pi.velocity.y());
        out += " 0\n";
    }
    out += "</DataArray>\n";
    arrayBegin("1", "Float64", "Pressure");
    for (const Particle& pi : particles) {
        appendG(out, pi.pressure);
        out += '\n';
    }
    out += "</DataArray>\n";
    arrayBegin("1", "Float64", "Number Density Ratio");
    for (const Particle& pi : particles) {
        appendG(out, pi.numberDensityRatio);
        out += '\n';
    }
    out += "</DataArray>\n";
    arrayBegin("1", "Int32", "Boundary Condition");
    for (const Particle& pi : particles) {
        appendInt(out, static_cast<int>(pi.boundaryCondition));
        out += '\n';
    }
    out += "</DataArray>\n</PointData>\n<Cells>\n";
    out += "<DataArray type='Int32' Name='connectivity' format='ascii'>\n";
    for (std::size_t i = 0; i < n; i++) {
        appendInt(out, (long long) i);
        out += '\n';
    }
    out += "</DataArray>\n<DataArray type='Int32' Name='offsets' format='ascii'>\n";
    for (std::size_t i = 0; i < n; i++) {
        appendInt(out, (long long) i + 1);
        out += '\n';
    }
    out += "</DataArray>\n<DataArray type='UInt8' Name='types' format='ascii'>\n";
    for (std::size_t i = 0; i < n; i++) out += "1\n";
    out += "</DataArray>\n</Cells>\n</Piece>\n</UnstructuredGrid>\n</VTKFile>\n";
    if (!writeAll(outFilePath, out)) mpslog::error("Could not open target vtu file: " + outFilePath.string());
}
