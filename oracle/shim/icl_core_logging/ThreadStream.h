// Shim (test infrastructure only): just enough of icl_core's ThreadStream for the
// reference's cuda_vectors.h / cuda_matrices.h stream operators to compile.
// Not a copy of icl_core; the real library (Boost + TinyXML) is absent on this image.
#ifndef GVX_SHIM_ICL_THREADSTREAM_H
#define GVX_SHIM_ICL_THREADSTREAM_H
#include <iostream>
#include <sstream>
#include <string>
namespace icl_core { namespace logging {
class ThreadStream {
public:
  template <class T> ThreadStream& operator<<(const T& v) { std::cerr << v; return *this; }
  ThreadStream& operator<<(ThreadStream& (*manip)(ThreadStream&)) { return manip(*this); }
};
inline ThreadStream& endl(ThreadStream& s) { std::cerr << std::endl; return s; }
}}
#endif
