// Minimal stand-in for glog (absent in this image) so the reference sources compile in place. Checker only.
#pragma once
#include <iostream>
#include <sstream>
#include <cstdlib>
struct NullLog { template<class T> NullLog& operator<<(const T&) { return *this; } NullLog& operator<<(std::ostream&(*)(std::ostream&)) { return *this; } };
struct FatalLog { std::ostringstream s; template<class T> FatalLog& operator<<(const T& v) { s << v; return *this; } ~FatalLog() { std::cerr << "FATAL: " << s.str() << std::endl; std::abort(); } };
#define LOG(sev) LOG_##sev
#define LOG_INFO NullLog()
#define LOG_WARNING NullLog()
#define LOG_ERROR NullLog()
#define LOG_FATAL FatalLog()
