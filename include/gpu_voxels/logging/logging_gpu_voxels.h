// gpu_voxels/logging/logging_gpu_voxels.h -- the reference logs through icl_core (absent here); the facade keeps the
// macro names its callers use and writes to std::cerr (GV/logging/logging_gpu_voxels.h).
#ifndef GPU_VOXELS_B200_LOGGING_H
#define GPU_VOXELS_B200_LOGGING_H

#include <iostream>

namespace gpu_voxels {
struct GvLogStream { const char* name; };
static const GvLogStream Gpu_voxels = {"GpuVoxels"};
static const GvLogStream Gpu_voxels_helpers = {"GpuVoxelsHelpers"};
static const GvLogStream VoxelmapLog = {"Voxelmap"};
static const GvLogStream VoxellistLog = {"Voxellist"};
static const GvLogStream RobotLog = {"Robot"};
static const char endl = '\n';
}  // namespace gpu_voxels

namespace icl_core { namespace logging {
// the command line, kept for the option stand-in of include/icl_core_config/Config.h (icl_core parses its registered
// options inside logging::initialize)
struct CommandLine { int argc; char** argv; CommandLine() : argc(0), argv(0) {} };
inline CommandLine& commandLine() { static CommandLine c; return c; }
inline void initialize(int& argc, char** argv) { commandLine().argc = argc; commandLine().argv = argv; }
inline void initialize() {}
inline void shutdown() {}
}}  // namespace icl_core::logging

#define GVX_LOG_(level, stream, msg) do { std::cerr << "<" level " " << (stream).name << "> " << msg; } while (0)
#define LOGGING_ERROR(stream, msg) GVX_LOG_("Error", stream, msg)
#define LOGGING_WARNING(stream, msg) GVX_LOG_("Warning", stream, msg)
#define LOGGING_INFO(stream, msg) GVX_LOG_("Info", stream, msg)
#define LOGGING_DEBUG(stream, msg) do { } while (0)
#define LOGGING_TRACE(stream, msg) do { } while (0)
#define LOGGING_ERROR_C(stream, cls, msg) GVX_LOG_("Error", stream, #cls ": " << msg)
#define LOGGING_WARNING_C(stream, cls, msg) GVX_LOG_("Warning", stream, #cls ": " << msg)
#define LOGGING_INFO_C(stream, cls, msg) GVX_LOG_("Info", stream, #cls ": " << msg)
#define LOGGING_DEBUG_C(stream, cls, msg) do { } while (0)
#define LOGGING_TRACE_C(stream, cls, msg) do { } while (0)
#endif
