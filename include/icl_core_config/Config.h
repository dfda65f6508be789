// icl_core_config/Config.h -- stand-in for the one piece of icl_core's option handling the reference's example programs
// use (examples/SweptVolumeVsEnvironment.cpp:68-73, RobotVsEnvironment.cpp:31): register a command-line option, read
// its value after icl_core::logging::initialize(argc, argv). icl_core is a separate library (not part of the
// reference tree, not installed here); this header exists so that those programs compile unchanged against the
// facade. An installation that has the real icl_core should put it first on the include path.
//   GetoptParameter("device-identifier:", "id", "help")   a trailing ':' = the option takes a value
//   accepted forms: --device-identifier=VALUE, --device-identifier VALUE, -id VALUE
#ifndef GPU_VOXELS_B200_ICL_CORE_CONFIG_CONFIG_H
#define GPU_VOXELS_B200_ICL_CORE_CONFIG_CONFIG_H

#include <string>
#include <vector>

#include <gpu_voxels/logging/logging_gpu_voxels.h>

namespace icl_core {
namespace config {

class GetoptParameter
{
public:
  GetoptParameter(const std::string& option, const std::string& short_option, const std::string& help,
                  bool is_prefix = false)
    : m_option(option), m_short_option(short_option), m_help(help), m_has_value(false), m_is_prefix(is_prefix)
  {
    if (!m_option.empty() && m_option[m_option.size() - 1] == ':')
    {
      m_option.erase(m_option.size() - 1);
      m_has_value = true;
    }
  }
  const std::string& option() const { return m_option; }
  const std::string& shortOption() const { return m_short_option; }
  const std::string& help() const { return m_help; }
  bool hasValue() const { return m_has_value; }
  bool isPrefixOption() const { return m_is_prefix; }

private:
  std::string m_option, m_short_option, m_help;
  bool m_has_value, m_is_prefix;
};

class Getopt
{
public:
  static Getopt& instance() { static Getopt g; return g; }
  void addParameter(const GetoptParameter& parameter) { m_parameters.push_back(parameter); }

  bool paramOptPresent(const std::string& name) const { std::string v; return find(name, v); }
  std::string paramOpt(const std::string& name) const { std::string v; find(name, v); return v; }

private:
  Getopt() {}
  // looks the option up in the command line icl_core::logging::initialize was given; "yes" for a present flag
  bool find(const std::string& name, std::string& value) const
  {
    const GetoptParameter* p = 0;
    for (size_t i = 0; i < m_parameters.size(); ++i)
      if (m_parameters[i].option() == name) p = &m_parameters[i];
    if (!p) return false;
    const icl_core::logging::CommandLine& cl = icl_core::logging::commandLine();
    const std::string long_form = "--" + p->option(), short_form = "-" + p->shortOption();
    for (int i = 1; i < cl.argc; ++i)
    {
      const std::string arg = cl.argv[i];
      if (p->hasValue() && arg.compare(0, long_form.size() + 1, long_form + "=") == 0)
      {
        value = arg.substr(long_form.size() + 1);
        return true;
      }
      if (arg == long_form || (!p->shortOption().empty() && arg == short_form))
      {
        if (!p->hasValue()) { value = "yes"; return true; }
        if (i + 1 < cl.argc) { value = cl.argv[i + 1]; return true; }
      }
    }
    return false;
  }
  std::vector<GetoptParameter> m_parameters;
};

inline void addParameter(const GetoptParameter& parameter) { Getopt::instance().addParameter(parameter); }

}  // namespace config
}  // namespace icl_core
#endif
