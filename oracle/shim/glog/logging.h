/* TEST INFRASTRUCTURE — stand-in for <glog/logging.h> (glog is not installed here): LOG(x) << ... is swallowed. */
#ifndef ORACLE_SHIM_GLOG_H
#define ORACLE_SHIM_GLOG_H
#include <ostream>
namespace oracle_shim
{
struct NullLog
{
    template <class T>
    NullLog& operator<<(const T&)
    {
        return *this;
    }
};
}  // namespace oracle_shim
#define LOG(level) oracle_shim::NullLog()
#define DLOG(level) oracle_shim::NullLog()
#endif
