// Push/pull microbenchmark ON the reference (our own program against the unmodified MiniPs library):
// the baseline of BASELINE.json config 5 ("keys/s and GB/s vs batch size ... against the reference ZMQ
// build").  One process per node (config file like the reference's), one worker and one server thread per
// node, a double table (the reference's KVPairs is fixed to double) range-partitioned evenly over the nodes, ASP so
// nothing gates.
//
// For each batch size n the worker times, with std::chrono around the reference's own calls,
//   push: Add(keys, vals) + Clock()  followed by a Get of one key per shard (Add is fire-and-forget in the
//         reference, so those replies mark the point where every shard has applied the push)
//   pull: Get(keys, &vals)
// for keys spread evenly over the whole key space ("spread": every shard receives n / nodes keys) and for
// one contiguous run inside a single remote shard ("contig").  It prints one line per case:
//   RESULT node=<id> mode=<spread|contig> n=<keys> push_ms=<median> pull_ms=<median>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <numeric>
#include <string>
#include <vector>

#include "base/node_utils.hpp"
#include "driver/engine.hpp"
#include "gflags/gflags.h"
#include "glog/logging.h"
#include "worker/kv_client_table.hpp"

DEFINE_int32(my_id, -1, "node id of this process");
DEFINE_string(config_file, "", "node list: id:host:port per line");
DEFINE_int32(max_keys, 10000000, "size of the key space");
DEFINE_int32(reps, 7, "timed repetitions per case");
DEFINE_string(sizes, "1000,10000,100000,1000000,10000000", "batch sizes");

namespace {
double median(std::vector<double> v) {
  std::sort(v.begin(), v.end());
  return v[v.size() / 2];
}
double ms_since(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}
}  // namespace

int main(int argc, char** argv) {
  gflags::ParseCommandLineFlags(&argc, &argv, true);
  google::InitGoogleLogging(argv[0]);
  using namespace minips;
  std::vector<Node> nodes = ParseFile(FLAGS_config_file);
  CHECK(CheckValidNodeIds(nodes));
  Node me = GetNodeById(nodes, FLAGS_my_id);
  std::vector<int> sizes;
  {
    std::string s = FLAGS_sizes;
    size_t pos = 0;
    while (pos < s.size()) {
      size_t c = s.find(',', pos);
      if (c == std::string::npos) c = s.size();
      sizes.push_back(std::atoi(s.substr(pos, c - pos).c_str()));
      pos = c + 1;
    }
  }

  Engine engine(me, nodes);
  engine.StartEverything();
  const uint64_t max_keys = static_cast<uint64_t>(FLAGS_max_keys);
  const uint64_t per = max_keys / nodes.size();
  std::vector<third_party::Range> ranges;
  for (size_t i = 0; i < nodes.size(); ++i)
    ranges.push_back({per * i, i + 1 == nodes.size() ? max_keys : per * (i + 1)});
  const uint32_t table_id = engine.CreateTable<double>(ranges, ModelType::ASP, StorageType::Vector, 0);
  engine.Barrier();

  MLTask task;
  std::vector<WorkerAlloc> alloc;
  for (auto& n : nodes) alloc.push_back({n.id, 1});
  task.SetWorkerAlloc(alloc);
  task.SetTables({table_id});
  const int reps = FLAGS_reps;
  const size_t my_pos = std::find_if(nodes.begin(), nodes.end(), [&](const Node& n) { return n.id == me.id; }) -
                        nodes.begin();
  task.SetLambda([&, table_id](const Info& info) {
    auto table = info.CreateKVClientTable<double>(table_id);
    third_party::SArray<Key> drain_keys(nodes.size());  // one key per shard: its reply follows the shard's adds
    for (size_t i = 0; i < nodes.size(); ++i) drain_keys[i] = static_cast<Key>(per * i);
    third_party::SArray<double> drain_vals;
    for (int mode = 0; mode < 2; ++mode) {
      for (int n : sizes) {
        if (static_cast<uint64_t>(n) > max_keys) continue;
        if (mode == 1 && static_cast<uint64_t>(n) > per) continue;  // a contiguous run must fit one shard
        third_party::SArray<Key> keys(n);
        if (mode == 0) {
          const double stride = static_cast<double>(max_keys) / n;
          for (int i = 0; i < n; ++i) keys[i] = static_cast<Key>(i * stride);
        } else {
          const uint64_t base = per * ((my_pos + 1) % nodes.size());  // the next node's shard
          for (int i = 0; i < n; ++i) keys[i] = static_cast<Key>(base + i);
        }
        third_party::SArray<double> vals(n, 0.5);
        third_party::SArray<double> out;
        std::vector<double> push_ms, pull_ms;
        for (int r = 0; r < reps + 2; ++r) {  // two untimed warm-up rounds
          auto t0 = std::chrono::steady_clock::now();
          table->Add(keys, vals);
          table->Clock();
          table->Get(drain_keys, &drain_vals);
          const double tp = ms_since(t0);
          t0 = std::chrono::steady_clock::now();
          table->Get(keys, &out);
          const double tg = ms_since(t0);
          CHECK_EQ(out.size(), keys.size());
          if (r >= 2) {
            push_ms.push_back(tp);
            pull_ms.push_back(tg);
          }
        }
        std::printf("RESULT node=%u mode=%s n=%d push_ms=%.4f pull_ms=%.4f\n", me.id, mode == 0 ? "spread"
            : "contig", n,
                    median(push_ms), median(pull_ms));
        std::fflush(stdout);
      }
    }
  });
  engine.Run(task);
  engine.StopEverything();
  return 0;
}
