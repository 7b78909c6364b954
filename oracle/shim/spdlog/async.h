// spdlog shim (see spdlog.h): async logging is not needed by the oracle build.
#pragma once
#include "spdlog.h"
