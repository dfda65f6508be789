// Shim (test infrastructure only): the LOGGING_* / DECLARE_LOG_STREAM macro names the
// reference's hot-path headers mention, mapped to stderr.
#ifndef GVX_SHIM_ICL_LOGGING_H
#define GVX_SHIM_ICL_LOGGING_H
#include "icl_core_logging/ThreadStream.h"
#define DECLARE_LOG_STREAM(name)
#define REGISTER_LOG_STREAM(name)
#define GVX_SHIM_LOG(msg) do { icl_core::logging::ThreadStream _s; _s << msg; } while (0)
#define LOGGING_ERROR(stream, msg)   GVX_SHIM_LOG(msg)
#define LOGGING_WARNING(stream, msg) GVX_SHIM_LOG(msg)
#define LOGGING_INFO(stream, msg)    GVX_SHIM_LOG(msg)
#define LOGGING_DEBUG(stream, msg)   do {} while (0)
#define LOGGING_TRACE(stream, msg)   do {} while (0)
#define LOGGING_ERROR_C(stream, cls, msg)   GVX_SHIM_LOG(msg)
#define LOGGING_WARNING_C(stream, cls, msg) GVX_SHIM_LOG(msg)
#define LOGGING_INFO_C(stream, cls, msg)    GVX_SHIM_LOG(msg)
#define LOGGING_DEBUG_C(stream, cls, msg)   do {} while (0)
#define LOGGING_TRACE_C(stream, cls, msg)   do {} while (0)
#endif
