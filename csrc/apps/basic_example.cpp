// basic_example.cpp — host-runtime smoke application (reference apps/basic/basic_example.cpp:25-76):
// 1000 keys range-partitioned over the nodes, SSP staleness 1, MapStorage, 10 workers per node,
// each doing 100 x (Get all keys, Add 0.5, Clock).
//   build/bin/basic_example --my_id=0 --config_file=config/localnodes
#include <iostream>

#include "core/driver.hpp"
#include "core/node_utils.hpp"

using namespace minips;  // NOLINT

int main(int argc, char** argv) {
  auto& flags = Flags::Get();
  flags.Define("my_id", "0", "node id of this process");
  flags.Define("config_file", "", "hostfile, one id:host:port per line");
  flags.Define("num_keys", "1000");
  flags.Define("num_workers_per_node", "10");
  flags.Define("num_iters", "100");
  flags.Parse(argc, argv);
  auto& ctx = Context::get_instance();
  std::vector<Node> nodes = ctx.get_string("config_file").empty()
                                ? std::vector<Node>{Node{0, "localhost", 14560}}
                                : ParseFile(ctx.get_string("config_file"));
  MINIPS_CHECK(CheckValidNodeIds(nodes) && CheckUniquePort(nodes));
  const Node my_node = GetNodeById(nodes, ctx.get_int32("my_id"));
  Engine engine(my_node, nodes);
  engine.StartEverything();

  const uint64_t n_keys = static_cast<uint64_t>(ctx.get_int32("num_keys"));
  std::vector<Range> ranges;
  const uint64_t per = n_keys / nodes.size();
  for (size_t i = 0; i < nodes.size(); ++i) ranges.emplace_back(per * i, i + 1 == nodes.size() ? n_keys
      : per * (i + 1));
  const uint32_t table = engine.CreateTable<double>(ranges, ModelType::SSP, StorageType::Map, 1);
  engine.Barrier();

  MLTask task;
  std::vector<WorkerAlloc> alloc;
  for (const auto& n : nodes) alloc.push_back({n.id, static_cast<uint32_t>(ctx.get_int32("num_workers_per_node"))});
  task.SetWorkerAlloc(alloc);
  task.SetTables({table});
  const int iters = ctx.get_int32("num_iters");
  task.SetLambda([&](const Info& info) {
    auto t = info.CreateKVClientTable<double>(table);
    std::vector<Key> keys(n_keys);
    for (uint64_t i = 0; i < n_keys; ++i) keys[i] = static_cast<Key>(i);
    std::vector<double> vals, half(n_keys, 0.5);
    for (int i = 0; i < iters; ++i) {
      t->Get(keys, &vals);
      t->Add(keys, half);
      t->Clock();
    }
  });
  engine.Run(task);

  // verification pass (after Run's closing barrier every add has been applied)
  bool ok = true;
  MLTask check;
  check.SetWorkerAlloc({{nodes.front().id, 1}});
  check.SetTables({table});
  const double expect = 0.5 * iters * ctx.get_int32("num_workers_per_node") * static_cast<double>(nodes.size());
  check.SetLambda([&](const Info& info) {
    auto t = info.CreateKVClientTable<double>(table);
    std::vector<Key> keys(n_keys);
    for (uint64_t i = 0; i < n_keys; ++i) keys[i] = static_cast<Key>(i);
    std::vector<double> vals;
    t->Get(keys, &vals);
    for (double v : vals) ok = ok && v == expect;
    std::cout << "BasicExample: every key == " << expect << ": " << (ok ? "true" : "false") << std::endl;
  });
  engine.Run(check);
  engine.StopEverything();
  return ok ? 0 : 1;
}
