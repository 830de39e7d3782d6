// kmeans.cpp — mini-batch k-means on the HOST runtime (reference apps/kmeans/kmeans.cpp:154-297,
// apps/kmeans/kmeans_helper.hpp).  Table 1: K*D centroid coordinates (row-major), table 2: K
// per-cluster counts.  Worker 0 seeds the centres (random | kmeans++ | kmeans_parallel) in an
// init task by adding them onto the zero table; then every worker iterates
//   Get both tables -> for each point of its batch: nearest centre, lr = alpha / ++n_c,
//   c -= lr (c - x)  -> Add (new - old) -> Clock both tables
// exactly like the reference's sequential per-point update.  Every `report_interval` iterations
// the report worker prints the SSE of a 50-point sample (kmeans_helper.hpp:236-266).
// Fixes w.r.t. the reference: the distance helper subtracts instead of
// overwriting, the sample index stays in range, and count ranges never come out empty.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <iostream>
#include <limits>
#include <random>

#include "core/driver.hpp"
#include "core/file_io.hpp"
#include "core/lib.hpp"
#include "core/node_utils.hpp"

using namespace minips;  // NOLINT
using Centers = std::vector<std::vector<double>>;

static double SqDist(const SVMItem& x, const std::vector<double>& c, double c_norm) {
  // ||x - c||^2 = ||c||^2 + sum_j (x_j^2 - 2 x_j c_j) over the non-zeros of x
  double d = c_norm;
  for (const auto& f : x.first) d += f.second * f.second - 2.0 * f.second * c[f.first];
  return d;
}
static double Norm2(const std::vector<double>& c) {
  double s = 0;
  for (double v : c) s += v * v;
  return s;
}
static std::pair<int, double> Nearest(const SVMItem& x, const Centers& c, const std::vector<double>& norms) {
  int best = 0;
  double bd = std::numeric_limits<double>::max();
  for (size_t k = 0; k < c.size(); ++k) {
    const double d = SqDist(x, c[k], norms[k]);
    if (d < bd) { bd = d; best = static_cast<int>(k); }
  }
  return {best, bd};
}
static std::vector<double> Dense(const SVMItem& x, int D) {
  std::vector<double> v(D, 0.0);
  for (const auto& f : x.first)
    if (f.first < D) v[f.first] = f.second;
  return v;
}

static Centers InitCenters(const std::string& mode, int K, int D, const std::vector<SVMItem>& data, std::mt19937& rng) {
  Centers c;
  std::uniform_int_distribution<size_t> pick(0, data.size() - 1);
  if (mode == "random") {
    std::vector<size_t> used;
    while (static_cast<int>(c.size()) < K) {
      size_t i = pick(rng);
      if (std::find(used.begin(), used.end(), i) != used.end() && used.size() < data.size()) continue;
      used.push_back(i);
      c.push_back(Dense(data[i], D));
    }
    return c;
  }
  // D^2 sampling; kmeans_parallel oversamples for a few rounds and then reduces with kmeans++ on the candidates
  auto d2_sample = [&](const std::vector<SVMItem>& pts, const std::vector<double>& w, int k) {
    Centers out{Dense(pts[rng() % pts.size()], D)};
    std::vector<double> d2(pts.size(), std::numeric_limits<double>::max());
    while (static_cast<int>(out.size()) < k) {
      const double n = Norm2(out.back());
      double total = 0;
      for (size_t i = 0; i < pts.size(); ++i) {
        d2[i] = std::min(d2[i], std::max(0.0, SqDist(pts[i], out.back(), n)));
        total += d2[i] * w[i];
      }
      double r = std::uniform_real_distribution<double>(0, total)(rng);
      size_t chosen = pts.size() - 1;
      for (size_t i = 0; i < pts.size(); ++i) {
        r -= d2[i] * w[i];
        if (r <= 0) { chosen = i; break; }
      }
      out.push_back(Dense(pts[chosen], D));
    }
    return out;
  };
  if (mode == "kmeans++") return d2_sample(data, std::vector<double>(data.size(), 1.0), K);
  // kmeans_parallel
  std::vector<SVMItem> cand{data[pick(rng)]};
  for (int round = 0; round < 5; ++round) {
    std::vector<double> d2(data.size());
    double total = 0;
    for (size_t i = 0; i < data.size(); ++i) {
      double best = std::numeric_limits<double>::max();
      for (const auto& cd : cand) {
        auto dc = Dense(cd, D);
        best = std::min(best, std::max(0.0, SqDist(data[i], dc, Norm2(dc))));
      }
      d2[i] = best;
      total += best;
    }
    for (size_t i = 0; i < data.size(); ++i)
      if (total > 0 && std::uniform_real_distribution<double>(0, 1)(rng) < std::min(1.0,
          2.0 * K * d2[i] / total)) cand.push_back(data[i]);
    if (cand.size() > static_cast<size_t>(20 * K)) break;
  }
  while (cand.size() < static_cast<size_t>(K)) cand.push_back(data[pick(rng)]);
  std::vector<double> w(cand.size(), 0.0);
  Centers cc;
  std::vector<double> cn;
  for (const auto& cd : cand) { cc.push_back(Dense(cd, D)); cn.push_back(Norm2(cc.back())); }
  for (const auto& x : data) w[Nearest(x, cc, cn).first] += 1.0;
  return d2_sample(cand, w, K);
}

int main(int argc, char** argv) {
  auto& f = Flags::Get();
  for (auto kv : std::vector<std::pair<const char*, const char*>>{
           {"my_id", "0"}, {"config_file", ""}, {"input", ""}, {"hdfs_namenode", "localhost"}, {"hdfs_namenode_port",
               "9000"},
           {"assigner_master_port", "19201"}, {"kModelType", "SSP"}, {"kStorageType", "Vector"}, {"num_dims", "123"},
           {"K", "2"}, {"batch_size", "100"}, {"num_iters", "10"}, {"kStaleness", "0"}, {"alpha", "0.1"},
           {"kmeans_init_mode", "random"}, {"report_interval", "2"}, {"report_worker", "0"}, {"report_prefix", ""},
           {"num_nodes", "0"}, {"num_workers_per_node", "1"}, {"num_servers_per_node",
               "1"}, {"num_local_load_thread", "2"},
           {"heartbeat_interval", "-1"}, {"block_size", "4194304"}, {"seed", "1"}})
    f.Define(kv.first, kv.second);
  f.Parse(argc, argv);
  auto& ctx = Context::get_instance();
  // validation (kmeans.cpp:59-70)
  MINIPS_CHECK(!ctx.get_string("config_file").empty()) << "--config_file is required";
  MINIPS_CHECK(ctx.get_int32("num_workers_per_node") > 0 && ctx.get_int32("num_workers_per_node") <= 20);
  MINIPS_CHECK(ctx.get_int32("num_servers_per_node") > 0 && ctx.get_int32("num_servers_per_node") <= 20);
  MINIPS_CHECK(ctx.get_int32("kStaleness") >= 0 && ctx.get_int32("kStaleness") <= 5);
  const std::string mode = ctx.get_string("kmeans_init_mode");
  MINIPS_CHECK(mode == "random" || mode == "kmeans++"
      || mode == "kmeans_parallel") << "unknown kmeans_init_mode " << mode;
  std::vector<Node> nodes = ParseFile(ctx.get_string("config_file"));
  const Node my_node = GetNodeById(nodes, ctx.get_int32("my_id"));
  const int K = ctx.get_int32("K"), D = ctx.get_int32("num_dims");

  std::vector<SVMItem> data;
  DataManager::Config cfg;
  cfg.master_host = nodes.front().hostname;
  cfg.master_port = ctx.get_int32("assigner_master_port");
  cfg.url = ctx.get_string("input");
  cfg.num_local_load_thread = ctx.get_int32("num_local_load_thread");
  cfg.block_size = static_cast<size_t>(ctx.get_int64("block_size"));
  lib::DataLoader<SVMItem, std::vector<SVMItem>>::load(
      cfg, my_node, nodes, [](const std::string& l, SVMItem* s) { return lib::Parser::parse_libsvm(l, s); }, &data);
  MINIPS_CHECK(!data.empty()) << "node " << my_node.id << " received no data";

  Engine engine(my_node, nodes);
  const int S = ctx.get_int32("num_servers_per_node");
  engine.StartEverything(S);
  const std::string mt = ctx.get_string("kModelType");
  const ModelType model = mt == "BSP" ? ModelType::BSP : (mt == "ASP" ? ModelType::ASP : ModelType::SSP);
  const size_t shards = nodes.size() * S;
  auto split = [&](uint64_t n) {
    std::vector<Range> r;
    for (size_t i = 0; i < shards; ++i) r.emplace_back(n * i / shards, n * (i + 1) / shards);  // never drops keys
    return r;
  };
  const uint32_t t_c = engine.CreateTable<double>(split(static_cast<uint64_t>(K) * D), model, StorageType::Vector,
      ctx.get_int32("kStaleness"));
  const uint32_t t_n = engine.CreateTable<double>(split(K), model, StorageType::Vector, ctx.get_int32("kStaleness"));
  engine.Barrier();

  std::vector<Key> keys_c(static_cast<size_t>(K) * D), keys_n(K);
  for (size_t i = 0; i < keys_c.size(); ++i) keys_c[i] = static_cast<Key>(i);
  for (int i = 0; i < K; ++i) keys_n[i] = static_cast<Key>(i);

  MLTask init;
  init.SetWorkerAlloc({{nodes.front().id, 1}});
  init.SetTables({t_c, t_n});
  init.SetLambda([&](const Info& info) {
    const auto t0 = std::chrono::steady_clock::now();
    std::mt19937 rng(ctx.get_int32("seed"));
    Centers c = InitCenters(mode, K, D, data, rng);
    std::vector<double> flat;
    for (auto& row : c) flat.insert(flat.end(), row.begin(), row.end());
    auto tc = info.CreateKVClientTable<double>(t_c);
    auto tn = info.CreateKVClientTable<double>(t_n);
    tc->Add(keys_c, flat);  // "set" = add onto zero (kmeans.cpp:185)
    tc->Clock();
    tn->Clock();
    std::cout << "[init] " << mode << " seeding took "
              << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() << " ms"
              << std::endl;
  });
  engine.Run(init);  // collective: nodes without a worker of the task still take part in its barriers
  engine.Barrier();

  const int W = ctx.get_int32("num_workers_per_node");
  const int per_worker = std::max(1, ctx.get_int32("batch_size") / W);
  const int iters = ctx.get_int32("num_iters");
  const double alpha = ctx.get_double("alpha");
  MLTask task;
  std::vector<WorkerAlloc> alloc;
  for (const auto& n : nodes) alloc.push_back({n.id, static_cast<uint32_t>(W)});
  task.SetWorkerAlloc(alloc);
  task.SetTables({t_c, t_n});
  task.SetLambda([&](const Info& info) {
    auto tc = info.CreateKVClientTable<double>(t_c);
    auto tn = info.CreateKVClientTable<double>(t_n);
    std::mt19937 rng(ctx.get_int32("seed") * 100 + info.worker_id);
    size_t startpt = rng() % data.size();
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<double> pull, members;
    for (int it = 0; it < iters; ++it) {
      tc->Get(keys_c, &pull);
      tn->Get(keys_n, &members);
      Centers cur(K, std::vector<double>(D));
      for (int k = 0; k < K; ++k) std::copy(pull.begin() + static_cast<size_t>(k) * D,
          pull.begin() + static_cast<size_t>(k + 1) * D, cur[k].begin());
      std::vector<double> norms(K), cnt0 = members;
      for (int k = 0; k < K; ++k) norms[k] = Norm2(cur[k]);
      for (int b = 0; b < per_worker; ++b) {
        const SVMItem& x = data[startpt];
        startpt = (startpt + 1) % data.size();
        const int c = Nearest(x, cur, norms).first;
        const double lr = alpha / ++members[c];
        std::vector<double> dx = Dense(x, D);
        for (int j = 0; j < D; ++j) cur[c][j] -= lr * (cur[c][j] - dx[j]);
        norms[c] = Norm2(cur[c]);
      }
      std::vector<double> delta(keys_c.size()), dn(K);
      for (int k = 0; k < K; ++k)
        for (int j = 0; j < D; ++j)
          delta[static_cast<size_t>(k) * D + j] = cur[k][j] - pull[static_cast<size_t>(k) * D + j];
      for (int k = 0; k < K; ++k) dn[k] = members[k] - cnt0[k];
      tc->Add(keys_c, delta);
      tc->Clock();
      tn->Add(keys_n, dn);
      tn->Clock();
      const int rep = ctx.get_int32("report_interval");
      if (rep > 0 && (it + 1) % rep == 0 && static_cast<int>(info.worker_id) == ctx.get_int32("report_worker")) {
        double sse = 0;
        const int samples = std::min<size_t>(50, data.size());
        for (int s = 0; s < samples; ++s) sse += std::max(0.0, Nearest(data[rng() % data.size()], cur, norms).second);
        sse *= static_cast<double>(data.size()) / samples;
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        std::cout << "iter " << it + 1 << "\tSSE " << sse << "\t" << ms << " ms" << std::endl;
        if (!ctx.get_string("report_prefix").empty()) {
          io::ofstream out(ctx.get_string("report_prefix") + "_" + ctx.get_string("my_id"),
              std::ios_base::out | std::ios_base::app);
          out << it + 1 << "\t" << sse << "\t" << ms << "\n";
        }
      }
    }
    std::cout << "Total training time: "
              << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count()
              << " ms on worker "
              << info.worker_id << std::endl;
  });
  engine.Run(task);
  engine.StopEverything();
  return 0;
}
