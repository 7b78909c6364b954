#include "settings.hpp"

#include "log.hpp"
#include "yaml_lite.hpp"

#include <algorithm>
#include <cstdlib>

void Settings::load(const std::filesystem::path& inputYamlPath) {
    try {
        const yaml_lite::Node root = yaml_lite::parseFile(inputYamlPath.string());

        dim              = root["dim"].asInt();
        particleDistance = root["particle distance"].asDouble();
        dt               = root["dt"].asDouble();
        finishTime       = root["finish time"].asDouble();
        outputInterval   = root["output interval"].asDouble();
        cflCondition     = root["CFL condition"].asDouble();

        const auto& rx = root["domain range X"];
        const auto& ry = root["domain range Y"];
        const auto& rz = root["domain range Z"];
        domain = Domain(rx[0].asDouble(), rx[1].asDouble(), ry[0].asDouble(), ry[1].asDouble(), rz[0].asDouble(), rz[1].asDouble());

        density            = root["density"].asDouble();
        kinematicViscosity = root["kinematic viscosity"].asDouble();

        const auto& g = root["gravity"];
        gravity       = Eigen::Vector3d(g[0].asDouble(), g[1].asDouble(), g[2].asDouble());

        // radius = ratio * l0: one rounded product each, as reference src/settings.cpp:45-53
        const auto& re                   = root["effective radius ratio"];
        effectiveRadius.pressure         = re["pressure"].asDouble() * particleDistance;
        effectiveRadius.viscosity        = re["viscosity"].asDouble() * particleDistance;
        effectiveRadius.surfaceDetection = re["surface detection"].asDouble() * particleDistance;
        effectiveRadius.max = std::max({effectiveRadius.pressure, effectiveRadius.viscosity, effectiveRadius.surfaceDetection});

        thresholdForSurfaceDetection = root["threshold for surface detection"].asDouble();

        const auto& ps          = root["pressure stabilization"];
        quasiCompressibility.on = ps["quasi compressibility"]["on"].asBool();
        if (quasiCompressibility.on) quasiCompressibility.compressibility = ps["quasi compressibility"]["compressibility"].asDouble();
        higherOrderSourceTerm.on = ps["higher order source term"]["on"].asBool();
        if (higherOrderSourceTerm.on) higherOrderSourceTerm.gamma = ps["higher order source term"]["gamma"].asDouble();
        relaxationCoefficientForPressure = root["relaxation coefficient for pressure"].asDouble();

        collisionDistance        = root["collision distance ratio"].asDouble() * particleDistance;
        coefficientOfRestitution = root["coefficient of restitution"].asDouble();

        inputCsvPath = root["inputCsvPath"].asString();
    } catch (const std::exception& e) {
        mpslog::error(std::string("Could not load setting file ") + inputYamlPath.string() + ": " + e.what());
        std::exit(EXIT_FAILURE);
    }
}
