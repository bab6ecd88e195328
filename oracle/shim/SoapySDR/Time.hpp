// oracle/shim/SoapySDR/Time.hpp -- TEST INFRASTRUCTURE (see Types.hpp)
#pragma once
namespace SoapySDR {
inline long long ticksToTimeNs(const long long ticks, const double rate) { return (long long)((double)ticks * 1e9 / rate); }
inline long long timeNsToTicks(const long long ns, const double rate) { return (long long)((double)ns * rate / 1e9); }
}  // namespace SoapySDR
