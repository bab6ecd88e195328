// oracle/shim/SoapySDR/Formats.h -- TEST INFRASTRUCTURE (see Types.hpp): the one stream format SoapySDDC offers.
#pragma once
#define SOAPY_SDR_CF32 "CF32"
