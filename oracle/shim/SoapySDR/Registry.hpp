// oracle/shim/SoapySDR/Registry.hpp -- TEST INFRASTRUCTURE (see Types.hpp): remembers the last registered factory.
#pragma once
#include <string>

#include "Device.hpp"

#define SOAPY_SDR_ABI_VERSION "0.8-shim"

namespace SoapySDR {
typedef KwargsList (*FindFunction)(const Kwargs&);
typedef Device* (*MakeFunction)(const Kwargs&);
class Registry {
public:
    Registry(const std::string& name, const FindFunction& find, const MakeFunction& make, const std::string& abi) {
        last_name() = name;
        last_make() = make;
        (void)find;
        (void)abi;
    }
    static std::string& last_name() {
        static std::string s;
        return s;
    }
    static MakeFunction& last_make() {
        static MakeFunction f = nullptr;
        return f;
    }
};
}  // namespace SoapySDR
