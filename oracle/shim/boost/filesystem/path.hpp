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
  // boost semantics as far as the reference needs them: "" / p == p; a separator is inserted when the
  // left side is non-empty and does not end in one; an absolute right side is appended as it is
  path operator/(const path& r) const {
    if (m_s.empty()) return r;
    if (r.m_s.empty()) return *this;
    if (m_s[m_s.size() - 1] == '/' || r.m_s[0] == '/') return path(m_s + r.m_s);
    return path(m_s + "/" + r.m_s);
  }
private:
  std::string m_s;
};
}}
#endif
