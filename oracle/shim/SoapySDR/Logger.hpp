// oracle/shim/SoapySDR/Logger.hpp -- TEST INFRASTRUCTURE (see Types.hpp): log calls go to stderr, SSI ("O" overflow
// marks) is counted so the driver can report it.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <string>

enum SoapySDRLogLevel { SOAPY_SDR_FATAL = 1, SOAPY_SDR_ERROR = 3, SOAPY_SDR_WARNING = 4, SOAPY_SDR_INFO = 6, SOAPY_SDR_DEBUG = 7, SOAPY_SDR_SSI = 9 };

inline int& soapy_shim_ssi_count() {
    static int n = 0;
    return n;
}
inline void SoapySDR_logf(const SoapySDRLogLevel level, const char* fmt, ...) {
    if (level > SOAPY_SDR_WARNING) return;
    va_list ap;
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    fputc('\n', stderr);
}
namespace SoapySDR {
inline void log(const SoapySDRLogLevel level, const std::string& msg) {
    if (level == SOAPY_SDR_SSI) soapy_shim_ssi_count()++;
    else if (level <= SOAPY_SDR_WARNING) fprintf(stderr, "%s\n", msg.c_str());
}
}  // namespace SoapySDR
