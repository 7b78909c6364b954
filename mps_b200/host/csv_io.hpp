// csv_io.hpp — reading the reference's particle CSV (layout of reference src/exporter.cpp:19-52 /
// inputs/*/input.csv: row 1 "Time (s),<t>", row 2 "Number of Particles,<n>", row 3 column names, then one
// row per particle) and the two float spellings its writer produces (reference csv-parser
// single_include/csv.hpp:6641-6681: std::to_string under clang, a 5-decimal digit loop under gcc).
#pragma once

#include <charconv>
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>

namespace csv_io {

// Split one CSV record (RFC 4180 quoting honoured; the reference's files never quote).
inline std::vector<std::string> splitRecord(const std::string& line) {
    std::vector<std::string> out;
    std::string cur;
    bool quoted = false;
    for (std::size_t i = 0; i < line.size(); i++) {
        char c = line[i];
        if (quoted) {
            if (c == '"' && i + 1 < line.size() && line[i + 1] == '"') {
                cur += '"';
                i++;
            } else if (c == '"') {
                quoted = false;
            } else {
                cur += c;
            }
        } else if (c == '"') {
            quoted = true;
        } else if (c == ',') {
            out.push_back(cur);
            cur.clear();
        } else if (c != '\r') {
            cur += c;
        }
    }
    out.push_back(cur);
    return out;
}

enum class FloatStyle {
    Fixed6,   // "%f": what a clang build of the reference writes (its shipped input.csv files)
    Trunc5Gcc // what a gcc build of the reference writes: 5 decimals, truncated, ".0" for integers
};

inline std::string formatFloat(double value, FloatStyle style) {
    if (style == FloatStyle::Fixed6) { // std::to_chars(fixed, 6) is specified to print what "%f" prints, several times faster
        char buf[400];
        auto r = std::to_chars(buf, buf + sizeof buf, value, std::chars_format::fixed, 6);
        return std::string(buf, r.ptr);
    }
    // digit-by-digit, integral and fractional part apart; each digit = trunc(fmod(part, 10^k) / 10^(k-1))
    auto p10 = [](int k) {
        double r = 1.0;
        for (int i = 0; i < k; i++) r *= 10.0;
        return r;
    };
    std::string s;
    double whole = 0.0;
    double frac  = std::fabs(std::modf(value, &whole));
    whole        = std::fabs(whole);
    if (value < 0) s += '-';
    if (whole == 0) {
        s += '0';
    } else {
        int nd = 0;
        for (double t = whole; t >= 1; t /= 10) nd++;
        for (int k = nd; k > 0; k--) s += (char) ('0' + (int) (std::fmod(whole, p10(k)) / p10(k - 1)));
    }
    s += '.';
    if (frac > 0) {
        frac *= p10(5);
        for (int k = 5; k > 0; k--) s += (char) ('0' + (int) (std::fmod(frac, p10(k)) / p10(k - 1)));
    } else {
        s += '0';
    }
    return s;
}

} // namespace csv_io
