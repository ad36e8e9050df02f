// TEST INFRASTRUCTURE ONLY. Stand-in for google-glog (an absent submodule of the reference), reached through
// saiga/core/util/easylogging++.h: the logging macros swallow their operands, the CHECK family does not evaluate.
#pragma once
#include <iostream>
struct SnkNullLog
{
    template <class T>
    SnkNullLog& operator<<(const T&)
    {
        return *this;
    }
    SnkNullLog& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
};
#define VLOG(x) SnkNullLog()
#define LOG(x) SnkNullLog()
#define LOG_IF(x, c) SnkNullLog()
#define DLOG(x) SnkNullLog()
#define CHECK(x) SnkNullLog()
#define CHECK_EQ(a, b) SnkNullLog()
#define CHECK_NE(a, b) SnkNullLog()
#define CHECK_LT(a, b) SnkNullLog()
#define CHECK_LE(a, b) SnkNullLog()
#define CHECK_GT(a, b) SnkNullLog()
#define CHECK_GE(a, b) SnkNullLog()
#define CHECK_NOTNULL(a) (a)
#define VLOG_IS_ON(x) false
