// oracle/shim/SoapySDR/Device.hpp -- TEST INFRASTRUCTURE (see Types.hpp): the base class SoapySDDC derives from, and the
// constants its stream calls use (values as SoapySDR/Constants.h and Errors.h publish them).
#pragma once
#include <stdexcept>

#include "Logger.hpp"
#include "Types.hpp"

#define SOAPY_SDR_TX 0
#define SOAPY_SDR_RX 1
#define SOAPY_SDR_MORE_FRAGMENTS (1 << 5)
#define SOAPY_SDR_TIMEOUT (-1)
#define SOAPY_SDR_OVERFLOW (-4)

namespace SoapySDR {
class Stream;
class Device {
public:
    virtual ~Device() {}
};
}  // namespace SoapySDR
