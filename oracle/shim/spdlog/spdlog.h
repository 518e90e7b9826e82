// Minimal stand-in for spdlog (the reference pins spdlog 3335c380 as an un-vendored
// submodule; its sources are absent from /root/reference). TEST INFRASTRUCTURE ONLY:
// lets the unmodified reference sources that include Source/Logger.h compile for the
// oracle build (oracle/_ref). All log macros evaluate nothing.
#pragma once
#include <string>
#include <sstream>
#include <iomanip>
#include <ctime>
#define SPDLOG_LEVEL_TRACE 0
#define SPDLOG_LEVEL_DEBUG 1
#define SPDLOG_LEVEL_INFO 2
#define SPDLOG_LEVEL_WARN 3
#define SPDLOG_LEVEL_ERROR 4
#ifndef SPDLOG_ACTIVE_LEVEL
#define SPDLOG_ACTIVE_LEVEL SPDLOG_LEVEL_INFO
#endif
#define SPDLOG_DEBUG(...) (void)0
#define SPDLOG_INFO(...) (void)0
#define SPDLOG_WARN(...) (void)0
#define SPDLOG_ERROR(...) (void)0
namespace spdlog {
namespace level { enum level_enum : int { trace = 0, debug, info, warn, err, critical, off }; }
inline void set_level(level::level_enum) {}
inline void set_pattern(const std::string&) {}
}
