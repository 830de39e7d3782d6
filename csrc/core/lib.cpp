#include "core/lib.hpp"

#include <iomanip>
#include <limits>

#include "core/checkpoint_io.hpp"
#include "core/context.hpp"
#include "core/file_io.hpp"
#include "core/node_utils.hpp"

namespace minips {
namespace lib {

bool Parser::parse_libsvm(const std::string& line, SVMItem* item) {
  item->first.clear();
  item->second = 0;
  const char* p = line.c_str();
  char* end = nullptr;
  while (*p == ' ' || *p == '\t') ++p;
  if (*p == '\0' || *p == '#') return false;
  item->second = std::strtod(p, &end);
  if (end == p) return false;
  p = end;
  while (true) {
    while (*p == ' ' || *p == '\t') ++p;
    if (*p == '\0' || *p == '\r' || *p == '\n' || *p == '#') break;
    const long idx = std::strtol(p, &end, 10);
    if (end == p || *end != ':') break;
    p = end + 1;
    const double val = std::strtod(p, &end);
    if (end == p) break;
    p = end;
    item->first.emplace_back(static_cast<int>(idx - 1), val);
  }
  return true;
}

void SVMDumper::DumpSVMData(const std::vector<SVMItem>& data, const std::string& path) {
  const std::string p = path.empty() ? checkpoint::Path("worker") : path;
  io::ofstream out(p, std::ios_base::out | std::ios_base::trunc);
  MINIPS_CHECK(out.is_open()) << "cannot write " << p;
  out << std::setprecision(std::numeric_limits<double>::max_digits10);
  for (const auto& item : data) {
    out << item.second;
    for (const auto& f : item.first) out << " " << (f.first + 1) << ":" << f.second;
    out << "\n";
  }
}

std::vector<SVMItem> SVMDumper::LoadSVMData(const std::string& path) {
  std::vector<SVMItem> data;
  auto& ctx = Context::get_instance();
  const std::string p =
      path.empty() ? ctx.get_string("checkpoint_raw_prefix") + "worker_" + ctx.get_string("my_id") : path;
  io::ifstream in(p);
  if (!in.is_open()) return data;
  std::string line;
  SVMItem item;
  while (std::getline(in, line))
    if (Parser::parse_libsvm(line, &item)) data.push_back(item);
  return data;
}

void SVMDumper::DumpConfigData(const std::map<int, int>& iter_map, const std::string& path) {
  checkpoint::DumpConfigData(iter_map, path);
}

std::map<int, int> SVMDumper::LoadConfigData(const std::string& path) { return checkpoint::LoadConfigData(path); }

void SVMDumper::DumpScaleFile(const Node& node, const std::string& path) {
  io::ofstream out(path, std::ios_base::out | std::ios_base::trunc);
  MINIPS_CHECK(out.is_open()) << "cannot write " << path;
  out << node.id << ":" << node.hostname << ":" << node.port << "\n";
}

Node SVMDumper::LoadScaleFile(const std::string& path) {
  auto nodes = ParseFile(path);
  MINIPS_CHECK(!nodes.empty()) << "empty scale file " << path;
  return nodes.front();
}

}  // namespace lib
}  // namespace minips
