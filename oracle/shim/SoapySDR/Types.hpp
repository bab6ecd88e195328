// oracle/shim/SoapySDR -- TEST INFRASTRUCTURE.  SoapySDR is not installed here; these headers declare just enough of its
// public C++ API (names and shapes as SoapySDR 0.8 documents them) for the reference's SoapySDDC/*.cpp to compile in
// place, so that its stream buffers (SoapySDDC/Settings.cpp:14-35, Streaming.cpp:110-182) can be driven by
// oracle/ref_soapy_driver.cpp.  No SoapySDR code is reproduced: the module loader, registry and converters do not exist.
#pragma once
#include <map>
#include <string>
#include <vector>

namespace SoapySDR {
typedef std::map<std::string, std::string> Kwargs;
typedef std::vector<Kwargs> KwargsList;

class Range {
public:
    Range() : lo_(0), hi_(0), step_(0) {}
    Range(double lo, double hi, double step = 0.0) : lo_(lo), hi_(hi), step_(step) {}
    double minimum() const { return lo_; }
    double maximum() const { return hi_; }
    double step() const { return step_; }

private:
    double lo_, hi_, step_;
};
typedef std::vector<Range> RangeList;

class ArgInfo {
public:
    enum Type { BOOL, INT, FLOAT, STRING };
    std::string key, value, name, description, units;
    Type type = STRING;
    Range range;
    std::vector<std::string> options, optionNames;
};
typedef std::vector<ArgInfo> ArgInfoList;
}  // namespace SoapySDR
