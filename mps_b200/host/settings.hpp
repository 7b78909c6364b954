// settings.hpp — host mirror of the reference's Settings / Domain (reference src/settings.hpp:7-86):
// the same public fields (including `double dim`, sic, :39) and `load`, reading the same YAML keys.
#pragma once

#include "vec3.hpp"

#include <filesystem>

class Domain {
public:
    struct {
        double min, max, length;
    } x, y, z;

    Domain() = default;
    Domain(
        const double& xMin, const double& xMax, const double& yMin, const double& yMax, const double& zMin,
        const double& zMax
    ) {
        x = {xMin, xMax, xMax - xMin};
        y = {yMin, yMax, yMax - yMin};
        z = {zMin, zMax, zMax - zMin};
    }
};

class Settings {
public:
    // computational conditions
    double dim;
    double particleDistance;
    double dt;
    double finishTime;
    double outputInterval;
    double cflCondition;

    // domain
    Domain domain;

    // physical properties
    double density;
    double kinematicViscosity;

    // gravity
    Eigen::Vector3d gravity;

    // effective radius (already multiplied by the particle distance)
    struct {
        double pressure;
        double viscosity;
        double surfaceDetection;
        double max;
    } effectiveRadius;

    // surface detection
    double thresholdForSurfaceDetection;

    // pressure calculation
    struct {
        bool on;
        double compressibility;
    } quasiCompressibility{false, 0.0};
    struct {
        bool on;
        double gamma;
    } higherOrderSourceTerm{false, 0.0};
    double relaxationCoefficientForPressure;

    // collision
    double collisionDistance; // already multiplied by the particle distance
    double coefficientOfRestitution;

    // io
    std::filesystem::path inputCsvPath;

    // Reads the keys of reference src/settings.cpp:20-84. The reference logs a parse error and carries on
    // with indeterminate fields; here an unreadable or incomplete file is logged and then fatal.
    void load(const std::filesystem::path& inputYamlPath);
};
