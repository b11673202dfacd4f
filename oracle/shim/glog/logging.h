// Test-infrastructure shim (oracle/): the few glog macros the reference's codebook_builder.cpp /
// codebook_encoder.h use, so that those files compile from where they lie under /root/reference without
// the glog dependency (fetched from the network by the reference's CMake). Failed checks abort.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <sstream>
namespace oracle_shim {
struct CheckSink {
    bool ok;
    const char* what;
    std::ostringstream os;
    CheckSink(bool ok_, const char* w) : ok(ok_), what(w) {}
    ~CheckSink() {
        if (!ok) {
            std::fprintf(stderr, "CHECK failed: %s %s\n", what, os.str().c_str());
            std::abort();
        }
    }
    template <typename T> CheckSink& operator<<(const T& v) { if (!ok) os << v; return *this; }
};
}  // namespace oracle_shim
#define CHECK(c) ::oracle_shim::CheckSink((c), #c)
#define CHECK_LE(a, b) ::oracle_shim::CheckSink((a) <= (b), #a " <= " #b)
#define CHECK_LT(a, b) ::oracle_shim::CheckSink((a) < (b), #a " < " #b)
#define CHECK_GE(a, b) ::oracle_shim::CheckSink((a) >= (b), #a " >= " #b)
#define CHECK_GT(a, b) ::oracle_shim::CheckSink((a) > (b), #a " > " #b)
#define CHECK_EQ(a, b) ::oracle_shim::CheckSink((a) == (b), #a " == " #b)
#define CHECK_NE(a, b) ::oracle_shim::CheckSink((a) != (b), #a " != " #b)
