// oracle/shim/SoapySDR/Types.h -- TEST INFRASTRUCTURE (see Types.hpp): the C header only has to exist.
#pragma once
