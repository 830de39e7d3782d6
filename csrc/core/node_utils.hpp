// node_utils.hpp — hostfile parsing and node-list checks.
// Reference: base/node_utils.hpp:15-54, .cpp:19-109.  Hostfile lines are `id:hostname:port`;
// the file may live on hdfs:// (opened through io::ifstream).
#pragma once

#include <string>
#include <vector>

#include "core/base.hpp"

namespace minips {

std::vector<Node> ParseFile(const std::string& filename);
std::vector<Node> ParseLines(const std::string& text);
// With heartbeat_interval > 0 the node with id 1 is taken OUT of `nodes` and returned as the
// master (it hosts no servers or workers); otherwise an empty Node (id 0, is_master=false).
Node SelectMaster(std::vector<Node>& nodes, int heartbeat_interval);
bool CheckValidNodeIds(const std::vector<Node>& nodes);   // ids < 1000 (thread-id layout)
bool CheckUniquePort(const std::vector<Node>& nodes);     // per hostname
bool CheckConsecutiveIds(const std::vector<Node>& nodes); // 0,1,2,...
bool HasNode(const std::vector<Node>& nodes, uint32_t id);
Node GetNodeById(const std::vector<Node>& nodes, int id); // throws if absent

}  // namespace minips
