// lr_example.cpp — sparse logistic regression on the HOST runtime (CPU plumbing configuration).
//
// Same application as the reference's LRExample (apps/logistic_regression/lr_example.cpp:97-431,
// flag names :20-56): load this node's libsvm samples through the block assigner, create one
// range-partitioned table, and run per worker   Get(batch keys) -> sigmoid/gradient -> Add -> Clock,
// with checkpoint every `checkpoint_interval` iterations, master heartbeat + rollback, ForceQuit
// of data-less nodes, optional report file and straggler injection, final accuracy.
// Deviations: deltas are zeroed each iteration, CheckPoint works under every
// consistency model, injected stragglers are OFF by default.
//
//   build/bin/lr_example --my_id=0 --config_file=config/localnodes --input=/tmp/a.libsvm \
//       --num_dims=1000 --kModelType=BSP --batch_size=100 --num_iters=100 --heartbeat_interval=-1
#include <atomic>
#include <chrono>
#include <cmath>
#include <iostream>
#include <mutex>
#include <random>
#include <sstream>

#include "core/checkpoint_io.hpp"
#include "core/driver.hpp"
#include "core/file_io.hpp"
#include "core/lib.hpp"
#include "core/master.hpp"
#include "core/node_utils.hpp"

using namespace minips;  // NOLINT

static void DefineFlags() {
  auto& f = Flags::Get();
  f.Define("my_id", "0");
  f.Define("config_file", "");
  f.Define("hdfs_namenode", "localhost");
  f.Define("hdfs_namenode_port", "9000");
  f.Define("assigner_master_port", "19201");
  f.Define("input", "", "libsvm file or directory (hdfs:// maps to $MINIPS_HDFS_ROOT)");
  f.Define("kModelType", "SSP", "ASP / SSP / BSP");
  f.Define("kStorageType", "Vector", "Map / Vector");
  f.Define("num_dims", "20958");
  f.Define("batch_size", "1");
  f.Define("num_iters", "1000");
  f.Define("kStaleness", "0");
  f.Define("kSpeculation", "5");
  f.Define("kSparseSSPRecorderType", "None");
  f.Define("num_workers_per_node", "2");
  f.Define("num_local_load_thread", "2");
  f.Define("num_servers_per_node", "1");
  f.Define("with_injected_straggler", "0", "probability that an iteration sleeps U(0,100) ms");
  f.Define("alpha", "0.1");
  f.Define("init_dump", "false");
  f.Define("use_weight_file", "false");
  f.Define("checkpoint_toggle", "false");
  f.Define("checkpoint_interval", "100");
  f.Define("weight_file_prefix", "");
  f.Define("checkpoint_file_prefix", "");
  f.Define("checkpoint_raw_prefix", "");
  f.Define("heartbeat_interval", "-1");
  f.Define("relaunch_cmd", "");
  f.Define("report_prefix", "");
  f.Define("report_interval", "-1");
  f.Define("scale", "false");
  f.Define("scale_name", "");
  f.Define("scale_port", "14563");
  f.Define("scale_file", "");
  f.Define("block_size", "4194304", "bytes per input block handed out by the assigner");
  f.Define("seed", "1");
}

static double TestAccuracy(const std::vector<double>& w, const std::vector<SVMItem>& data) {
  if (data.empty()) return 0;
  size_t correct = 0;
  for (const auto& s : data) {
    double z = 0;
    for (const auto& f : s.first)
      if (static_cast<size_t>(f.first) < w.size()) z += w[f.first] * f.second;
    const double pred = 1.0 / (1.0 + std::exp(-z));
    const int y = s.second > 0 ? 1 : 0;
    correct += ((pred > 0.5) ? 1 : 0) == y;
  }
  return static_cast<double>(correct) / data.size();
}

static int Training(const Node& my_node, const std::vector<Node>& nodes, const Node& master, const Node& scale_node) {
  auto& ctx = Context::get_instance();
  const bool resume = ctx.get_bool("use_weight_file");
  if (resume) CheckFaultTolerance(3);

  // ---- data
  std::vector<SVMItem> data;
  if (resume) {
    data = lib::SVMDumper::LoadSVMData();
  } else {
    DataManager::Config cfg;
    cfg.master_host = nodes.front().hostname;
    cfg.master_port = ctx.get_int32("assigner_master_port");
    cfg.worker_host = my_node.hostname;
    cfg.hdfs_namenode = ctx.get_string("hdfs_namenode");
    cfg.hdfs_namenode_port = ctx.get_int32("hdfs_namenode_port");
    cfg.url = ctx.get_string("input");
    cfg.num_local_load_thread = ctx.get_int32("num_local_load_thread");
    cfg.block_size = static_cast<size_t>(ctx.get_int64("block_size"));
    lib::DataLoader<SVMItem, std::vector<SVMItem>>::load(
        cfg, my_node, nodes, [](const std::string& l, SVMItem* s) { return lib::Parser::parse_libsvm(l, s); }, &data);
  }
  MINIPS_LOG(0) << "node " << my_node.id << " loaded " << data.size() << " samples";

  Engine engine(my_node, nodes, master, scale_node);
  engine.StartEverything(ctx.get_int32("num_servers_per_node"));
  if (data.empty()) {  // lr_example.cpp:145-152
    engine.Barrier();
    engine.ForceQuit();
    engine.GetMailbox()->Stop(false);
    engine.StopSender();
    return 0;
  }
  if (!resume) engine.Barrier();

  const std::string mt = ctx.get_string("kModelType");
  const ModelType model = mt == "BSP" ? ModelType::BSP : (mt == "ASP" ? ModelType::ASP : ModelType::SSP);
  const StorageType storage = ctx.get_string("kStorageType") == "Map" ? StorageType::Map : StorageType::Vector;
  const uint32_t table = engine.CreateTable<double>(engine.getRanges(), model, storage, ctx.get_int32("kStaleness"));
  engine.SetDumpCallback([] {});
  if (ctx.get_bool("init_dump") && ctx.get_bool("checkpoint_toggle") && !resume) lib::SVMDumper::DumpSVMData(data);
  if (resume) {
    Context::get_instance().update_iteration(lib::SVMDumper::LoadConfigData());
    engine.RollBackServer();
    CheckFaultTolerance(4);
  }
  if (!resume) engine.Barrier();

  const int num_dims = ctx.get_int32("num_dims");
  const int batch = ctx.get_int32("batch_size");
  const int iters = ctx.get_int32("num_iters");
  const double alpha = ctx.get_double("alpha");
  const int ck_interval = std::max(1, ctx.get_int32("checkpoint_interval"));
  const bool ck = ctx.get_bool("checkpoint_toggle");
  const int report_interval = ctx.get_int32("report_interval");
  const double p_straggle = ctx.get_double("with_injected_straggler");

  MLTask task;
  std::vector<WorkerAlloc> alloc;
  for (const auto& n : engine.GetNodes()) alloc.push_back({n.id,
      static_cast<uint32_t>(ctx.get_int32("num_workers_per_node"))});
  task.SetWorkerAlloc(alloc);
  task.SetTables({table});
  std::mutex print_mu;
  std::atomic<bool> recovered_logged{false};
  task.SetLambda([&](const Info& info) {
    lib::BatchDataSampler<SVMItem> sampler(data, batch,
        static_cast<uint64_t>(ctx.get_int32("seed")) * 1000 + info.worker_id + 1);
    // all batches up front, like lr_example.cpp:231-236
    std::vector<SArray<Key>> future_keys;
    std::vector<std::vector<const SVMItem*>> future_ptrs;
    for (int i = 0; i < iters + ctx.get_int32("kSpeculation"); ++i) {
      sampler.random_start_point();
      future_keys.push_back(sampler.prepare_next_batch());
      future_ptrs.push_back(sampler.get_data_ptrs());
    }
    auto t = info.CreateKVClientTable<double>(table);
    SArray<double> params, deltas;
    std::mt19937 rng(info.worker_id + 17);
    const auto t0 = std::chrono::steady_clock::now();
    auto elapsed_ms = [&] { return std::chrono::duration<double,
        std::milli>(std::chrono::steady_clock::now() - t0).count(); };
    for (int i = Context::get_instance().get_iteration(info.worker_id); i < iters; ++i) {
      const SArray<Key>& keys = future_keys[i];
      if (keys.empty()) {
        t->Clock();
        continue;
      }
      t->Get(keys, &params);
      // the first Get of a relaunched node was answered: every node is back (lr_example.cpp:265-268)
      if (resume && !recovered_logged.exchange(true)) CheckFaultTolerance(5);
      if (engine.IsNeedRollBack()) {  // lr_example.cpp:265-288
        if (engine.IncRollBackCount() == 1) {
          engine.Barrier();  // meets the restarted node's fault barrier
          engine.RecoverEnd();
        } else {
          engine.WaitRecover();
        }
        i = Context::get_instance().get_iteration(info.worker_id) - 1;
        continue;
      }
      deltas.clear();
      deltas.resize(keys.size(), 0.0);
      for (const SVMItem* s : future_ptrs[i]) {
        double z = 0;
        size_t pos = 0;
        for (const auto& f : s->first) {  // merge-walk: sample indices and batch keys are both ascending
          while (pos < keys.size() && keys[pos] < static_cast<Key>(f.first)) ++pos;
          if (pos < keys.size() && keys[pos] == static_cast<Key>(f.first)) z += params[pos] * f.second;
        }
        const double pred = 1.0 / (1.0 + std::exp(-z));
        const double y = s->second > 0 ? 1.0 : 0.0;
        pos = 0;
        for (const auto& f : s->first) {
          while (pos < keys.size() && keys[pos] < static_cast<Key>(f.first)) ++pos;
          if (pos < keys.size() && keys[pos] == static_cast<Key>(f.first)) deltas[pos] += alpha * f.second * (y - pred);
        }
      }
      SArray<double> push;  // own copy: the send is asynchronous
      push.CopyFrom(deltas);
      t->Add(keys, push);
      t->Clock();
      Context::get_instance().set_iteration(info.worker_id, i + 1);
      if (ck && info.worker_id == 0 && (i + 1) % ck_interval == 0) {
        const double c0 = elapsed_ms();
        t->CheckPoint();
        MINIPS_LOG(0) << "checkpoint at iteration " << i + 1 << " took " << elapsed_ms() - c0 << " ms";
      }
      if (report_interval > 0 && (i + 1) % report_interval == 0 && info.worker_id == 0
          && !ctx.get_string("report_prefix").empty()) {
        io::ofstream out(ctx.get_string("report_prefix") + "_" + ctx.get_string("my_id"),
            std::ios_base::out | std::ios_base::app);
        out << i + 1 << "\t" << elapsed_ms() << "\n";
      }
      if (p_straggle > 0 && std::uniform_real_distribution<double>(0, 1)(rng) < p_straggle)
        std::this_thread::sleep_for(std::chrono::milliseconds(rng() % 100));
    }
    const double total = elapsed_ms();
    if (info.thread_id % 1000 == 100) {  // first worker of every node (lr_example.cpp:360-365)
      std::vector<Key> all(num_dims);
      for (int k = 0; k < num_dims; ++k) all[k] = static_cast<Key>(k);
      std::vector<double> w;
      t->Get(all, &w);
      std::ostringstream line;
      line << "[node " << my_node.id << "] accuracy " << TestAccuracy(w, data) << " on " << data.size()
           << " local samples\n";
      std::lock_guard<std::mutex> lk(print_mu);  // worker threads share std::cout: one line, one write
      std::cout << line.str() << std::flush;
    }
    std::ostringstream line;
    line << "Total training time: " << total << " ms on worker " << info.worker_id << "\n";
    std::lock_guard<std::mutex> lk(print_mu);
    std::cout << line.str() << std::flush;
  });
  engine.Run(task);
  engine.StopEverything();
  return 0;
}

int main(int argc, char** argv) {
  DefineFlags();
  Flags::Get().Parse(argc, argv);
  auto& ctx = Context::get_instance();
  MINIPS_CHECK(!ctx.get_string("config_file").empty()) << "--config_file is required";
  std::vector<Node> nodes = ParseFile(ctx.get_string("config_file"));
  Node master = SelectMaster(nodes, ctx.get_int32("heartbeat_interval"));
  MINIPS_CHECK(CheckValidNodeIds(nodes) && CheckUniquePort(nodes));
  Node scale_node;
  if (ctx.get_bool("scale") && !ctx.get_string("scale_name").empty()) {
    scale_node = Node{static_cast<uint32_t>(ctx.get_int32("my_id")), ctx.get_string("scale_name"),
        ctx.get_int32("scale_port")};
    if (!ctx.get_string("scale_file").empty()) lib::SVMDumper::DumpScaleFile(scale_node, ctx.get_string("scale_file"));
    ctx.set("has_scale_node", true);
    ctx.set("scale_node_id", static_cast<int32_t>(scale_node.id));
  }
  const int my_id = ctx.get_int32("my_id");
  if (master.is_master && static_cast<int>(master.id) == my_id) {
    Master m(master, nodes);
    m.StartMaster();
    m.StopMaster();
    return 0;
  }
  MINIPS_CHECK(HasNode(nodes, my_id)) << "node " << my_id << " is not in " << ctx.get_string("config_file");
  return Training(GetNodeById(nodes, my_id), nodes, master, scale_node);
}
