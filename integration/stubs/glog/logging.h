// Stand-in for <glog/logging.h>: the log statements compile and go nowhere.
#ifndef B200_STUB_GLOG_LOGGING_H
#define B200_STUB_GLOG_LOGGING_H
#include <ostream>
struct b200_null_log
{
    template <class T>
    b200_null_log& operator<<(const T&)
    {
        return *this;
    }
};
#define LOG(severity) b200_null_log()
#define DLOG(severity) b200_null_log()
#endif
