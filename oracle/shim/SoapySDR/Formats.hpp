// oracle/shim/SoapySDR/Formats.hpp -- TEST INFRASTRUCTURE (see Types.hpp)
#pragma once
#include "Formats.h"
