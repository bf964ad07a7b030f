// spdlog stand-in: the reference's detectors only log (distributed_cv_feature.cpp:1216) -- logging is a no-op here.
#pragma once
namespace spdlog
{
template <typename... A> inline void info(A&&...) {}
template <typename... A> inline void debug(A&&...) {}
template <typename... A> inline void warn(A&&...) {}
template <typename... A> inline void error(A&&...) {}
} // namespace spdlog
