#pragma once
#include <string>
namespace lhb {
extern thread_local std::string g_host_err;
int host_fail(const char* fmt, ...);
}  // namespace lhb
