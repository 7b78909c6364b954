#include "exporter.hpp"

#include "log.hpp"

#include <algorithm>
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {
// operator<<(ostream, double) with default flags == "%g" with precision 6 (reference src/exporter.cpp:83-131)
inline void appendG(std::string& s, double v) { // std::to_chars(general, 6) == "%g"
    char buf[40];
    auto r = std::to_chars(buf, buf + sizeof buf, v, std::chars_format::general, 6);
    s.append(buf, (std::size_t) (r.ptr - buf));
}
inline void appendInt(std::string& s, long long v) {
    char buf[24];
    auto r = std::to_chars(buf, buf + sizeof buf, v);
    s.append(buf, (std::size_t) (r.ptr - buf));
}
// Formats [0, n) in parallel: `body(out, begin, end)` appends the text of particles [begin, end) to `out`; the
// pieces are concatenated in index order, so the file is the one a serial loop would write.
template <class Body>
void appendParallel(std::string& out, std::size_t n, Body body) {
    const std::size_t minChunk = 1u << 16;
    std::size_t workers = std::min<std::size_t>(std::max(1u, std::thread::hardware_concurrency()), (n + minChunk - 1) / minChunk);
    if (workers <= 1) {
        body(out, (std::size_t) 0, n);
        return;
    }
    std::vector<std::string> piece(workers);
    std::vector<std::thread> pool;
    const std::size_t per = (n + workers - 1) / workers;
    for (std::size_t w = 0; w < workers; w++)
        pool.emplace_back([&, w] {
            const std::size_t b = w * per, e = std::min(n, b + per);
            if (b < e) {
                piece[w].reserve((e - b) * 96);
                body(piece[w], b, e);
            }
        });
    for (auto& t : pool) t.join();
    for (auto& p : piece) out += p;
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
    const csv_io::FloatStyle style = floatStyle;
    appendParallel(out, particles.size(), [&particles, style](std::string& o, std::size_t b, std::size_t e) {
        for (std::size_t i = b; i < e; i++) {
            const Particle& pi = particles[i];
            if (pi.type == ParticleType::Ghost) continue;
            appendInt(o, pi.id);
            o += ',';
            appendInt(o, static_cast<int>(pi.type));
            const double v[8] = {pi.position.x(), pi.position.y(), pi.position.z(), pi.velocity.x(), pi.velocity.y(), pi.velocity.z(),
                                 pi.pressure, pi.numberDensityRatio};
            for (double x : v) {
                o += ',';
                o += csv_io::formatFloat(x, style);
            }
            o += '\n';
        }
    });
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
    appendParallel(out, n, [&particles](std::string& o, std::size_t b, std::size_t e) {
        for (std::size_t i = b; i < e; i++) {
            const Particle& pi = particles[i];
            appendG(o, pi.position.x()); // ghosts stay in the VTU (SURVEY B-7)
            o += ' ';
            appendG(o, pi.position.y());
            o += ' ';
            appendG(o, pi.position.z());
            o += '\n';
        }
    });
    out += "</DataArray>\n</Points>\n<PointData>\n";
    arrayBegin("1", "Int32", "Type");
    appendParallel(out, n, [&particles](std::string& o, std::size_t b, std::size_t e) {
        for (std::size_t i = b; i < e; i++) {
            const Particle& pi = particles[i];
            appendInt(o, static_cast<int>(pi.type));
            o += '\n';
        }
    });
    out += "</DataArray>\n";
    arrayBegin("3", "Float64", "Velocity");
    appendParallel(out, n, [&particles](std::string& o, std::size_t b, std::size_t e) {
        for (std::size_t i = b; i < e; i++) {
            const Particle& pi = particles[i];
            appendG(o, pi.velocity.x()); // z component written as a literal zero, as the reference
            o += ' ';
            appendG(o, pi.velocity.y());
            o += " 0\n";
        }
    });
    out += "</DataArray>\n";
    arrayBegin("1", "Float64", "Pressure");
    appendParallel(out, n, [&particles](std::string& o, std::size_t b, std::size_t e) {
        for (std::size_t i = b; i < e; i++) {
            const Particle& pi = particles[i];
            appendG(o, pi.pressure);
            o += '\n';
        }
    });
    out += "</DataArray>\n";
    arrayBegin("1", "Float64", "Number Density Ratio");
    appendParallel(out, n, [&particles](std::string& o, std::size_t b, std::size_t e) {
        for (std::size_t i = b; i < e; i++) {
            const Particle& pi = particles[i];
            appendG(o, pi.numberDensityRatio);
            o += '\n';
        }
    });
    out += "</DataArray>\n";
    arrayBegin("1", "Int32", "Boundary Condition");
    appendParallel(out, n, [&particles](std::string& o, std::size_t b, std::size_t e) {
        for (std::size_t i = b; i < e; i++) {
            const Particle& pi = particles[i];
            appendInt(o, static_cast<int>(pi.boundaryCondition));
            o += '\n';
        }
    });
    out += "</DataArray>\n</PointData>\n<Cells>\n";
    out += "<DataArray type='Int32' Name='connectivity' format='ascii'>\n";
    appendParallel(out, n, [](std::string& o, std::size_t b, std::size_t e) {
        for (std::size_t i = b; i < e; i++) {
            appendInt(o, (long long) i);
            o += '\n';
        }
    });
    out += "</DataArray>\n<DataArray type='Int32' Name='offsets' format='ascii'>\n";
    appendParallel(out, n, [](std::string& o, std::size_t b, std::size_t e) {
        for (std::size_t i = b; i < e; i++) {
            appendInt(o, (long long) i + 1);
            o += '\n';
        }
    });
    out += "</DataArray>\n<DataArray type='UInt8' Name='types' format='ascii'>\n";
    for (std::size_t i = 0; i < n; i++) out += "1\n";
    out += "</DataArray>\n</Cells>\n</Piece>\n</UnstructuredGrid>\n</VTKFile>\n";
    if (!writeAll(outFilePath, out)) mpslog::error("Could not open target vtu file: " + outFilePath.string());
}
