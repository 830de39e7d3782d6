#include "core/context.hpp"

#include <cstdlib>

namespace minips {

Flags& Flags::Get() {
  static Flags f;
  return f;
}

void Flags::Define(const std::string& name, const std::string& default_value, const std::string& help) {
  std::lock_guard<std::mutex> lk(mu_);
  if (!values_.count(name)) values_[name] = default_value;
  help_[name] = help;
}

std::vector<std::string> Flags::Parse(int argc, char** argv) {
  std::vector<std::string> args;
  for (int i = 1; i < argc; ++i) args.emplace_back(argv[i]);
  return Parse(args);
}

std::vector<std::string> Flags::Parse(const std::vector<std::string>& args) {
  std::vector<std::string> positional;
  for (size_t i = 0; i < args.size(); ++i) {
    const std::string& a = args[i];
    if (a.size() < 2 || a[0] != '-') {
      positional.push_back(a);
      continue;
    }
    std::string body = a.substr(a[1] == '-' ? 2 : 1);
    std::string name, value;
    const size_t eq = body.find('=');
    if (eq != std::string::npos) {
      name = body.substr(0, eq);
      value = body.substr(eq + 1);
    } else {
      name = body;
      bool known_bool = false;
      {
        std::lock_guard<std::mutex> lk(mu_);
        auto it = values_.find(name);
        known_bool = it != values_.end() && (it->second == "true" || it->second == "false");
        if (!known_bool && name.rfind("no", 0) == 0) {
          auto it2 = values_.find(name.substr(2));
          if (it2 != values_.end() && (it2->second == "true" || it2->second == "false")) {
            name = name.substr(2);
            value = "false";
            known_bool = true;
          }
        } else if (known_bool) {
          value = "true";
        }
      }
      if (!known_bool) {
        if (i + 1 < args.size() && (args[i + 1].empty() || args[i + 1][0] != '-')) value = args[++i];
        else value = "true";
      }
    }
    Set(name, value);
  }
  return positional;
}

void Flags::Set(const std::string& name, const std::string& value) {
  std::lock_guard<std::mutex> lk(mu_);
  std::string v = value;
  if (v == "True" || v == "TRUE") v = "true";
  if (v == "False" || v == "FALSE") v = "false";
  values_[name] = v;
}

bool Flags::Has(const std::string& name) const {
  std::lock_guard<std::mutex> lk(mu_);
  return values_.count(name) > 0;
}

std::string Flags::GetString(const std::string& name) const {
  std::lock_guard<std::mutex> lk(mu_);
  auto it = values_.find(name);
  return it == values_.end() ? std::string() : it->second;
}

std::map<std::string, std::string> Flags::All() const {
  std::lock_guard<std::mutex> lk(mu_);
  return values_;
}

std::string Flags::Usage() const {
  std::lock_guard<std::mutex> lk(mu_);
  std::ostringstream os;
  for (const auto& kv : values_) {
    os << "  --" << kv.first << "=" << kv.second;
    auto h = help_.find(kv.first);
    if (h != help_.end() && !h->second.empty()) os << "    " << h->second;
    os << "\n";
  }
  return os.str();
}

void Flags::Reset() {
  std::lock_guard<std::mutex> lk(mu_);
  values_.clear();
  help_.clear();
}

// ------------------------------------------------------------------ Context
Context& Context::get_instance() {
  static Context c;
  return c;
}

std::string Context::get_string(const std::string& key) const {
  {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = overrides_.find(key);
    if (it != overrides_.end()) return it->second;
  }
  return Flags::Get().GetString(key);
}
int32_t Context::get_int32(const std::string& key) const { return std::atoi(get_string(key).c_str()); }
int64_t Context::get_int64(const std::string& key) const { return std::atoll(get_string(key).c_str()); }
double Context::get_double(const std::string& key) const { return std::atof(get_string(key).c_str()); }
bool Context::get_bool(const std::string& key) const {
  const std::string v = get_string(key);
  return v == "true" || v == "True" || v == "1";
}

void Context::set(const std::string& key, const std::string& value) {
  std::lock_guard<std::mutex> lk(mu_);
  overrides_[key] = value;
}
void Context::set(const std::string& key, int32_t value) { set(key, std::to_string(value)); }
void Context::set(const std::string& key, int64_t value) { set(key, std::to_string(value)); }
void Context::set(const std::string& key, double value) { set(key, std::to_string(value)); }
void Context::set(const std::string& key, bool value) { set(key, std::string(value ? "true" : "false")); }

int Context::normalize_worker(int worker_id) const {
  // under --scale the reference folds worker ids onto [0, num_workers_per_node) (context.hpp:45-47)
  if (get_bool("scale")) {
    const int per = get_int32("num_workers_per_node");
    if (per > 0) return worker_id % per;
  }
  return worker_id;
}
int Context::get_iteration(int worker_id) const {
  const int w = normalize_worker(worker_id);
  std::lock_guard<std::mutex> lk(mu_);
  auto it = iteration_.find(w);
  return it == iteration_.end() ? 0 : it->second;
}
void Context::set_iteration(int worker_id, int iter) {
  const int w = normalize_worker(worker_id);
  std::lock_guard<std::mutex> lk(mu_);
  iteration_[w] = iter;
}
std::map<int, int> Context::get_iteration_map() const {
  std::lock_guard<std::mutex> lk(mu_);
  return iteration_;
}
void Context::update_iteration(const std::map<int, int>& m) {
  std::lock_guard<std::mutex> lk(mu_);
  for (const auto& kv : m) iteration_[kv.first] = kv.second;
}
void Context::clear() {
  std::lock_guard<std::mutex> lk(mu_);
  overrides_.clear();
  iteration_.clear();
}

}  // namespace minips
