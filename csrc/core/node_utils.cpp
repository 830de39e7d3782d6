#include "core/node_utils.hpp"

#include <set>
#include <sstream>

#include "core/file_io.hpp"

namespace minips {

std::vector<Node> ParseLines(const std::string& text) {
  std::vector<Node> nodes;
  std::istringstream in(text);
  std::string line;
  while (std::getline(in, line)) {
    while (!line.empty() && (line.back() == '\r' || line.back() == ' ' || line.back() == '\t')) line.pop_back();
    size_t s = line.find_first_not_of(" \t");
    if (s == std::string::npos || line[s] == '#') continue;
    line = line.substr(s);
    const size_t a = line.find(':');
    const size_t b = line.rfind(':');
    MINIPS_CHECK(a != std::string::npos && b != a) << "bad hostfile line '" << line << "' (want id:host:port)";
    Node n;
    n.id = static_cast<uint32_t>(std::stoul(line.substr(0, a)));
    n.hostname = line.substr(a + 1, b - a - 1);
    n.port = std::stoi(line.substr(b + 1));
    nodes.push_back(n);
  }
  return nodes;
}

std::vector<Node> ParseFile(const std::string& filename) {
  io::ifstream in(filename);
  MINIPS_CHECK(in.is_open()) << "cannot open hostfile " << filename;
  std::stringstream ss;
  ss << in.rdbuf();
  return ParseLines(ss.str());
}

Node SelectMaster(std::vector<Node>& nodes, int heartbeat_interval) {
  Node master;
  if (heartbeat_interval <= 0) return master;
  for (auto it = nodes.begin(); it != nodes.end(); ++it) {
    if (it->id == 1) {
      master = *it;
      master.is_master = true;
      nodes.erase(it);
      break;
    }
  }
  return master;
}

bool CheckValidNodeIds(const std::vector<Node>& nodes) {
  for (const auto& n : nodes)
    if (n.id >= 1000) return false;
  return true;
}

bool CheckUniquePort(const std::vector<Node>& nodes) {
  std::set<std::pair<std::string, int>> seen;
  for (const auto& n : nodes)
    if (!seen.emplace(n.hostname, n.port).second) return false;
  return true;
}

bool CheckConsecutiveIds(const std::vector<Node>& nodes) {
  for (size_t i = 0; i < nodes.size(); ++i)
    if (nodes[i].id != i) return false;
  return true;
}

bool HasNode(const std::vector<Node>& nodes, uint32_t id) {
  for (const auto& n : nodes)
    if (n.id == id) return true;
  return false;
}

Node GetNodeById(const std::vector<Node>& nodes, int id) {
  for (const auto& n : nodes)
    if (static_cast<int>(n.id) == id) return n;
  MINIPS_CHECK(false) << "Node " << id << " is not in the given node list";
  return Node();
}

}  // namespace minips
