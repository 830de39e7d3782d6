// context.hpp — flag registry + the Context singleton.
//
// The reference mirrors every gflag into a string map at first use so that library code can
// read application flags by name (base/context.hpp:14-63, base/context.cpp:13-61), and keeps
// the per-worker iteration map that checkpoint/rollback persist (context.hpp:36-53).  gflags
// is not available here, so `Flags` is a small registry with the same command-line syntax
// (--name=value, --name value, --noname for booleans).
#pragma once

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "core/base.hpp"

namespace minips {

class Flags {
 public:
  static Flags& Get();
  void Define(const std::string& name, const std::string& default_value, const std::string& help = "");
  // Parses argv; unknown --flags are kept (library code may read them).  Returns positional args.
  std::vector<std::string> Parse(int argc, char** argv);
  std::vector<std::string> Parse(const std::vector<std::string>& args);
  void Set(const std::string& name, const std::string& value);
  bool Has(const std::string& name) const;
  std::string GetString(const std::string& name) const;
  std::map<std::string, std::string> All() const;
  std::string Usage() const;
  void Reset();  // tests

 private:
  mutable std::mutex mu_;
  std::map<std::string, std::string> values_;
  std::map<std::string, std::string> help_;
};

class Context {
 public:
  static Context& get_instance();

  // Missing keys read as "" / 0 / false, like the reference (context.cpp:36-43).
  int32_t get_int32(const std::string& key) const;
  int64_t get_int64(const std::string& key) const;
  double get_double(const std::string& key) const;
  bool get_bool(const std::string& key) const;
  std::string get_string(const std::string& key) const;

  void set(const std::string& key, const std::string& value);
  void set(const std::string& key, int32_t value);
  void set(const std::string& key, int64_t value);
  void set(const std::string& key, double value);
  void set(const std::string& key, bool value);
  void set(const std::string& key, const char* value) { set(key, std::string(value)); }

  // per-worker iteration map (what SVMDumper::DumpConfigData persists)
  int get_iteration(int worker_id) const;
  void set_iteration(int worker_id, int iter);
  std::map<int, int> get_iteration_map() const;
  void update_iteration(const std::map<int, int>& m);

  void clear();  // tests

 private:
  Context() = default;
  int normalize_worker(int worker_id) const;
  mutable std::mutex mu_;
  std::map<std::string, std::string> overrides_;
  std::map<int, int> iteration_;
};

}  // namespace minips
