// TEST INFRASTRUCTURE ONLY.  Stand-in for <spdlog/logger.h> when the unmodified reference sources are compiled for oracle/_ref:
// the image's spdlog headers (1.14) do not build against its fmt (12), and the reference's genotype sinks only log through
// `gelex::logging::get()->error(...)`.  Messages are dropped; the exceptions the reference throws next to them are what is tested.
#pragma once
#include <string>

namespace spdlog {
class logger {
   public:
    template <typename... Args> void trace(Args&&...) {}
    template <typename... Args> void debug(Args&&...) {}
    template <typename... Args> void info(Args&&...) {}
    template <typename... Args> void warn(Args&&...) {}
    template <typename... Args> void error(Args&&...) {}
    template <typename... Args> void critical(Args&&...) {}
};
}  // namespace spdlog
