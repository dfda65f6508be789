// Shim (test infrastructure only): boost::filesystem::path as used by one inline helper
// in the reference's common_defines.h. Boost is absent on this image.
#ifndef GVX_SHIM_BOOST_FS_PATH_HPP
#define GVX_SHIM_BOOST_FS_PATH_HPP
#include <string>
namespace boost { namespace filesystem {
class path {
public:
  path() {}
  path(const char* s) : m_s(s) {}
  path(const std::string& s) : m_s(s) {}
  const std::string& string() const { return m_s; }
private:
  std::string m_s;
};
}}
#endif
