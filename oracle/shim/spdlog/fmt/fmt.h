// See ../spdlog.h. fmt::format is only reached through the LOGE macro, which the shim
// turns into a no-op, so a declaration-free namespace is enough.
#pragma once
#include <string>
namespace fmt {
template <typename... A> inline std::string format(const A&...) { return std::string(); }
}
