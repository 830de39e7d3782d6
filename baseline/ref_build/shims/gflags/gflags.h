// gflags/gflags.h shim — DEFINE_* / DECLARE_* / ParseCommandLineFlags / GetAllFlags.
#ifndef MINIPS_SHIM_GFLAGS_H_
#define MINIPS_SHIM_GFLAGS_H_
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <map>
#include <string>
#include <vector>

namespace gflags {
typedef int32_t int32;
typedef int64_t int64;
typedef uint64_t uint64;
struct CommandLineFlagInfo {
  std::string name, type, description, current_value, default_value, filename;
  bool has_validator_fn = false, is_default = true;
};
struct FlagEntry {
  std::string name, type, help, def;
  std::function<void(const std::string&)> set;
  std::function<std::string()> get;
};
inline std::map<std::string, FlagEntry>& Registry() {
  static std::map<std::string, FlagEntry> r;
  return r;
}
inline bool ParseBool(const std::string& v) {
  return v == "true" || v == "True" || v == "1" || v == "t" || v == "yes" || v.empty();
}
struct Registrar {
  Registrar(const char* name, const char* type, const char* help, bool* p, bool d) {
    Registry()[name] = {name, type, help, d ? "true" : "false", [p](const std::string& v) { *p = ParseBool(v); },
                        [p] { return std::string(*p ? "true" : "false"); }};
  }
  Registrar(const char* name, const char* type, const char* help, int32_t* p, int32_t d) {
    Registry()[name] = {name, type, help, std::to_string(d), [p](const std::string& v) { *p = std::atoi(v.c_str()); },
                        [p] { return std::to_string(*p); }};
  }
  Registrar(const char* name, const char* type, const char* help, uint32_t* p, uint32_t d) {
    Registry()[name] = {name, type, help, std::to_string(d),
                        [p](const std::string& v) { *p = static_cast<uint32_t>(std::strtoul(v.c_str(), nullptr, 10)); },
                        [p] { return std::to_string(*p); }};
  }
  Registrar(const char* name, const char* type, const char* help, int64_t* p, int64_t d) {
    Registry()[name] = {name, type, help, std::to_string(d), [p](const std::string& v) { *p = std::atoll(v.c_str()); },
                        [p] { return std::to_string(*p); }};
  }
  Registrar(const char* name, const char* type, const char* help, uint64_t* p, uint64_t d) {
    Registry()[name] = {name, type, help, std::to_string(d),
                        [p](const std::string& v) { *p = std::strtoull(v.c_str(), nullptr, 10); },
                        [p] { return std::to_string(*p); }};
  }
  Registrar(const char* name, const char* type, const char* help, double* p, double d) {
    Registry()[name] = {name, type, help, std::to_string(d), [p](const std::string& v) { *p = std::atof(v.c_str()); },
                        [p] { return std::to_string(*p); }};
  }
  Registrar(const char* name, const char* type, const char* help, std::string* p, const std::string& d) {
    Registry()[name] = {name, type, help, d, [p](const std::string& v) { *p = v; }, [p] { return *p; }};
  }
};
inline void GetAllFlags(std::vector<CommandLineFlagInfo>* out) {
  out->clear();
  for (auto& kv : Registry()) {
    CommandLineFlagInfo i;
    i.name = kv.second.name;
    i.type = kv.second.type;
    i.description = kv.second.help;
    i.default_value = kv.second.def;
    i.current_value = kv.second.get();
    i.is_default = i.current_value == i.default_value;
    out->push_back(i);
  }
}
inline void SetUsageMessage(const std::string&) {}
inline uint32_t ParseCommandLineFlags(int* argc, char*** argv, bool /*remove_flags*/) {
  for (int i = 1; i < *argc; ++i) {
    std::string a = (*argv)[i];
    if (a.size() < 2 || a[0] != '-') continue;
    a = a.substr(a[1] == '-' ? 2 : 1);
    std::string name = a, value;
    bool has_value = false;
    const size_t eq = a.find('=');
    if (eq != std::string::npos) {
      name = a.substr(0, eq);
      value = a.substr(eq + 1);
      has_value = true;
    }
    auto it = Registry().find(name);
    if (it == Registry().end()) {
      if (name.rfind("no", 0) == 0 && Registry().count(name.substr(2)) && Registry()[name.substr(2)].type == "bool") {
        Registry()[name.substr(2)].set("false");
        continue;
      }
      std::cerr << "ERROR: unknown command line flag '" << name << "'\n";
      std::exit(1);
    }
    if (!has_value) {
      if (it->second.type == "bool") value = "true";
      else if (i + 1 < *argc) value = (*argv)[++i];
    }
    it->second.set(value);
  }
  return 1;
}
}  // namespace gflags
namespace google {
using gflags::CommandLineFlagInfo;
using gflags::GetAllFlags;
using gflags::ParseCommandLineFlags;
using gflags::SetUsageMessage;
}  // namespace google

#define MINIPS_DEFINE_FLAG(ctype, tname, name, def, help)                                      \
  ctype FLAGS_##name = def;                                                                  \
  static ::gflags::Registrar minips_flag_reg_##name(#name, tname, help, &FLAGS_##name, def)
#define DEFINE_bool(name, def, help) MINIPS_DEFINE_FLAG(bool, "bool", name, def, help)
#define DEFINE_int32(name, def, help) MINIPS_DEFINE_FLAG(int32_t, "int32", name, def, help)
#define DEFINE_uint32(name, def, help) MINIPS_DEFINE_FLAG(uint32_t, "uint32", name, def, help)
#define DECLARE_uint32(name) extern uint32_t FLAGS_##name
#define DEFINE_int64(name, def, help) MINIPS_DEFINE_FLAG(int64_t, "int64", name, def, help)
#define DEFINE_uint64(name, def, help) MINIPS_DEFINE_FLAG(uint64_t, "uint64", name, def, help)
#define DEFINE_double(name, def, help) MINIPS_DEFINE_FLAG(double, "double", name, def, help)
#define DEFINE_string(name, def, help) \
  std::string FLAGS_##name = def;      \
  static ::gflags::Registrar minips_flag_reg_##name(#name, "string", help, &FLAGS_##name, std::string(def))
#define DECLARE_bool(name) extern bool FLAGS_##name
#define DECLARE_int32(name) extern int32_t FLAGS_##name
#define DECLARE_int64(name) extern int64_t FLAGS_##name
#define DECLARE_uint64(name) extern uint64_t FLAGS_##name
#define DECLARE_double(name) extern double FLAGS_##name
#define DECLARE_string(name) extern std::string FLAGS_##name
#endif
