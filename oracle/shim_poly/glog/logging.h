// oracle/shim_poly/glog/logging.h — TEST INFRASTRUCTURE.  Stand-in for the glog macros the reference's
// external/poly_traj sources use (glog is not installed in this image): CHECK* abort with a message on
// failure, LOG / VLOG swallow their stream.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <ostream>
#include <sstream>

namespace shim_glog {
struct Sink {
    template <typename T>
    Sink &operator<<(const T &) { return *this; }
    Sink &operator<<(std::ostream &(*)(std::ostream &)) { return *this; }
};
struct Fatal {
    const char *what;
    std::ostringstream os;
    explicit Fatal(const char *w) : what(w) {}
    template <typename T>
    Fatal &operator<<(const T &v)
    {
        os << v;
        return *this;
    }
    ~Fatal()
    {
        std::fprintf(stderr, "CHECK failed: %s %s\n", what, os.str().c_str());
        std::abort();
    }
};
template <typename T>
T &&check_not_null(T &&p, const char *what)
{
    if (p == nullptr) Fatal(what) << "is null";
    return static_cast<T &&>(p);
}
} // namespace shim_glog

#define LOG(severity) ::shim_glog::Sink()
#define VLOG(level) ::shim_glog::Sink()
#define DLOG(severity) ::shim_glog::Sink()
#define CHECK(cond) \
    if (cond) {     \
    } else          \
        ::shim_glog::Fatal(#cond)
#define CHECK_EQ(a, b) CHECK((a) == (b))
#define CHECK_NE(a, b) CHECK((a) != (b))
#define CHECK_LT(a, b) CHECK((a) < (b))
#define CHECK_LE(a, b) CHECK((a) <= (b))
#define CHECK_GT(a, b) CHECK((a) > (b))
#define CHECK_GE(a, b) CHECK((a) >= (b))
#define CHECK_NOTNULL(p) ::shim_glog::check_not_null((p), #p)
