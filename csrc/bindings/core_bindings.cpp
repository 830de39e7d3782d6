// core_bindings.cpp — pybind11 module `_minips_core`: the C++ host runtime from Python.
// Gives Python users the reference's C++ driver API on the CPU backend (Engine / MLTask / Info /
// KVClientTable<double>) plus the building blocks the tests and tools use directly.
#include <pybind11/functional.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include "core/bin_stream.hpp"
#include "core/checkpoint_io.hpp"
#include "core/driver.hpp"
#include "core/file_io.hpp"
#include "core/lib.hpp"
#include "core/master.hpp"
#include "core/node_utils.hpp"

namespace py = pybind11;
using namespace minips;  // NOLINT

namespace {

using Table = KVClientTable<double>;

std::vector<std::pair<int, std::vector<Key>>> SliceKeys(const RangePartitionManager& pm, const std::vector<Key>& keys) {
  std::vector<std::pair<int, SArray<Key>>> sliced;
  pm.Slice(SArray<Key>(keys), &sliced);
  std::vector<std::pair<int, std::vector<Key>>> out;
  for (auto& s : sliced) out.emplace_back(s.first, SArrayToVector(s.second));
  return out;
}

std::vector<std::tuple<int, std::vector<Key>, std::vector<double>>> SliceKVs(const RangePartitionManager& pm,
                                                                             const std::vector<Key>& keys,
                                                                             const std::vector<double>& vals) {
  std::vector<std::pair<int, AbstractPartitionManager::KVPairs>> sliced;
  pm.Slice(std::make_pair(SArray<Key>(keys), SArray<double>(vals)), &sliced);
  std::vector<std::tuple<int, std::vector<Key>, std::vector<double>>> out;
  for (auto& s : sliced) out.emplace_back(s.first, SArrayToVector(s.second.first), SArrayToVector(s.second.second));
  return out;
}

}  // namespace

PYBIND11_MODULE(_minips_core, m) {
  m.doc() = "minips_b200 C++ host runtime (actors, mailbox, storages, BSP/SSP/ASP, engine)";
  py::register_exception<CheckError>(m, "CheckError");
  m.def("set_log_level", &SetLogLevel);

  py::class_<Node>(m, "Node")
      .def(py::init([](uint32_t id, const std::string& host, int port, bool is_master) {
             Node n{id, host, port, is_master};
             return n;
           }),
           py::arg("id") = 0, py::arg("hostname") = "localhost", py::arg("port") = 0, py::arg("is_master") = false)
      .def_readwrite("id", &Node::id)
      .def_readwrite("hostname", &Node::hostname)
      .def_readwrite("port", &Node::port)
      .def_readwrite("is_master", &Node::is_master)
      .def("debug_string", &Node::DebugString);

  py::class_<Range>(m, "Range")
      .def(py::init<uint64_t, uint64_t>())
      .def("begin", &Range::begin)
      .def("end", &Range::end)
      .def("size", &Range::size);

  py::enum_<ModelType>(m, "ModelType").value("SSP", ModelType::SSP).value("BSP", ModelType::BSP).value("ASP",
      ModelType::ASP);
  py::enum_<StorageType>(m, "StorageType").value("Map", StorageType::Map).value("Vector", StorageType::Vector);

  // ---- node utils / flags / context
  m.def("parse_file", &ParseFile);
  m.def("parse_lines", &ParseLines);
  m.def("select_master", [](std::vector<Node> nodes, int hb) {
    Node master = SelectMaster(nodes, hb);
    return std::make_pair(master, nodes);
  });
  m.def("check_valid_node_ids", &CheckValidNodeIds);
  m.def("check_unique_port", &CheckUniquePort);
  m.def("check_consecutive_ids", &CheckConsecutiveIds);
  m.def("has_node", &HasNode);
  m.def("get_node_by_id", &GetNodeById);
  m.def("check_fault_tolerance", &CheckFaultTolerance);
  m.def("resolve_path", &io::ResolvePath);

  m.def("define_flag", [](const std::string& n, const std::string& d, const std::string& h) { Flags::Get().Define(n,
      d, h); },
        py::arg("name"), py::arg("default"), py::arg("help") = "");
  m.def("parse_flags", [](const std::vector<std::string>& args) { return Flags::Get().Parse(args); });
  m.def("all_flags", [] { return Flags::Get().All(); });
  m.def("context_set", [](const std::string& k, const std::string& v) { Context::get_instance().set(k, v); });
  m.def("context_get", [](const std::string& k) { return Context::get_instance().get_string(k); });
  m.def("context_clear", [] {
    Context::get_instance().clear();
    Flags::Get().Reset();
  });
  m.def("get_iteration", [](int w) { return Context::get_instance().get_iteration(w); });
  m.def("set_iteration", [](int w, int it) { Context::get_instance().set_iteration(w, it); });

  // ---- partitioning / ids / specs
  py::class_<RangePartitionManager>(m, "RangePartitionManager")
      .def(py::init([](const std::vector<uint32_t>& tids, const std::vector<std::pair<uint64_t, uint64_t>>& ranges) {
        std::vector<Range> r;
        for (auto& p : ranges) r.emplace_back(p.first, p.second);
        return new RangePartitionManager(tids, r);
      }))
      .def("get_num_servers", &RangePartitionManager::GetNumServers)
      .def("get_server_thread_ids", &RangePartitionManager::GetServerThreadIds)
      .def("slice_keys", &SliceKeys)
      .def("slice_kvs", &SliceKVs);

  py::class_<SimpleIdMapper>(m, "SimpleIdMapper")
      .def(py::init<const Node&, const std::vector<Node>&>())
      .def("init", &SimpleIdMapper::Init, py::arg("num_server_threads_per_node") = 1)
      .def("get_node_id_for_thread", &SimpleIdMapper::GetNodeIdForThread)
      .def("allocate_worker_thread", &SimpleIdMapper::AllocateWorkerThread)
      .def("deallocate_worker_thread", &SimpleIdMapper::DeallocateWorkerThread)
      .def("get_server_threads_for_id", &SimpleIdMapper::GetServerThreadsForId)
      .def("get_worker_helper_threads_for_id", &SimpleIdMapper::GetWorkerHelperThreadsForId)
      .def("get_worker_threads_for_id", &SimpleIdMapper::GetWorkerThreadsForId)
      .def("get_all_server_threads", &SimpleIdMapper::GetAllServerThreads);

  py::class_<WorkerAlloc>(m, "WorkerAlloc")
      .def(py::init([](uint32_t node, uint32_t n) { return WorkerAlloc{node, n}; }))
      .def_readwrite("node_id", &WorkerAlloc::node_id)
      .def_readwrite("num_workers", &WorkerAlloc::num_workers);

  py::class_<WorkerSpec>(m, "WorkerSpec")
      .def(py::init<const std::vector<WorkerAlloc>&>())
      .def("has_local_workers", &WorkerSpec::HasLocalWorkers)
      .def("get_local_workers", &WorkerSpec::GetLocalWorkers)
      .def("get_local_threads", &WorkerSpec::GetLocalThreads)
      .def("get_all_thread_ids", &WorkerSpec::GetAllThreadIds)
      .def("insert_worker_id_thread_id", &WorkerSpec::InsertWorkerIdThreadId)
      .def("get_num_workers", &WorkerSpec::GetNumWorkers);

  py::class_<ProgressTracker>(m, "ProgressTracker")
      .def(py::init<>())
      .def("init", &ProgressTracker::Init)
      .def("advance_and_get_changed_min_clock", &ProgressTracker::AdvanceAndGetChangedMinClock)
      .def("get_num_threads", &ProgressTracker::GetNumThreads)
      .def("get_progress", &ProgressTracker::GetProgress)
      .def("get_min_clock", &ProgressTracker::GetMinClock)
      .def("is_unique_min", &ProgressTracker::IsUniqueMin)
      .def("check_thread_valid", &ProgressTracker::CheckThreadValid)
      .def("delete_node", &ProgressTracker::DeleteNode)
      .def("dump_to", &ProgressTracker::DumpTo)
      .def("restore_from", &ProgressTracker::RestoreFrom);

  // ---- lib
  m.def("parse_libsvm", [](const std::string& line) -> py::object {
    SVMItem item;
    if (!lib::Parser::parse_libsvm(line, &item)) return py::none();
    return py::cast(std::make_pair(item.first, item.second));
  });
  m.def("load_config_data", [](const std::string& p) { return lib::SVMDumper::LoadConfigData(p); },
      py::arg("path") = "");
  m.def("dump_config_data", [](const std::map<int, int>& m2,
      const std::string& p) { lib::SVMDumper::DumpConfigData(m2, p); },
        py::arg("iter_map"), py::arg("path") = "");
  m.def("dump_scale_file", [](const Node& n, const std::string& p) { lib::SVMDumper::DumpScaleFile(n, p); },
        py::arg("node"), py::arg("path"));
  m.def("load_scale_file", [](const std::string& p) { return lib::SVMDumper::LoadScaleFile(p); }, py::arg("path"));

  // ---- engine
  py::class_<Table>(m, "KVClientTable")
      .def("get",
           [](Table& t, const std::vector<Key>& keys) {
             std::vector<double> vals;
             {
               py::gil_scoped_release rel;
               t.Get(keys, &vals);
             }
             return vals;
           })
      .def("add", [](Table& t, const std::vector<Key>& keys, const std::vector<double>& vals) { t.Add(keys, vals); })
      .def("clock", &Table::Clock)
      .def("checkpoint", &Table::CheckPoint, py::call_guard<py::gil_scoped_release>());

  py::class_<Info>(m, "Info")
      .def_readonly("thread_id", &Info::thread_id)
      .def_readonly("worker_id", &Info::worker_id)
      .def("create_kv_client_table", [](const Info& info,
          uint32_t table_id) { return info.CreateKVClientTable<double>(table_id); })
      .def("debug_string", &Info::DebugString);

  py::class_<MLTask>(m, "MLTask")
      .def(py::init<>())
      .def("set_lambda",
           [](MLTask& t, py::function fn) {
             // worker threads are C++ threads: take the GIL around the Python callable
             t.SetLambda([fn](const Info& info) {
               py::gil_scoped_acquire acq;
               fn(info);
             });
           })
      .def("set_worker_alloc", &MLTask::SetWorkerAlloc)
      .def("set_tables", &MLTask::SetTables)
      .def("is_setup", &MLTask::IsSetup);

  py::class_<Engine>(m, "Engine")
      .def(py::init([](const Node& node, const std::vector<Node>& nodes,
          const Node& master) { return new Engine(node, nodes, master); }),
           py::arg("node"), py::arg("nodes"), py::arg("master_node") = Node())
      .def("start_everything", &Engine::StartEverything, py::arg("num_server_threads_per_node") = 1,
           py::call_guard<py::gil_scoped_release>())
      .def("stop_everything", &Engine::StopEverything, py::call_guard<py::gil_scoped_release>())
      .def("create_table",
           [](Engine& e, const std::vector<std::pair<uint64_t, uint64_t>>& ranges, ModelType mt, StorageType st,
               int staleness) {
             std::vector<Range> r;
             for (auto& p : ranges) r.emplace_back(p.first, p.second);
             return e.CreateTable<double>(r, mt, st, staleness);
           },
           py::arg("ranges"), py::arg("model_type") = ModelType::SSP, py::arg("storage_type") = StorageType::Map,
           py::arg("staleness") = 0)
      .def("barrier", &Engine::Barrier, py::call_guard<py::gil_scoped_release>())
      .def("run", &Engine::Run, py::call_guard<py::gil_scoped_release>())
      .def("force_quit", &Engine::ForceQuit)
      .def("is_need_rollback", &Engine::IsNeedRollBack)
      .def("get_nodes", &Engine::GetNodes);

  py::class_<Master>(m, "Master")
      .def(py::init<const Node&, const std::vector<Node>&>())
      .def("start_master", &Master::StartMaster, py::call_guard<py::gil_scoped_release>())
      .def("stop_master", &Master::StopMaster, py::call_guard<py::gil_scoped_release>())
      .def("shutdown", &Master::Shutdown, py::call_guard<py::gil_scoped_release>())
      .def("rollback", &Master::RollBack, py::call_guard<py::gil_scoped_release>())
      .def("set_rollback_hook",
           [](Master& m2, py::function fn) {
             // runs on the master thread: take the GIL for the duration of the Python callback
             auto shared = std::make_shared<py::function>(std::move(fn));
             m2.SetRollBackHook([shared](uint32_t failed) {
               py::gil_scoped_acquire gil;
               (*shared)(failed);
             });
           })
      .def("set_scale_hook",
           [](Master& m2, py::function fn) {
             auto shared = std::make_shared<py::function>(std::move(fn));
             m2.SetScaleHook([shared](uint32_t node_id) {
               py::gil_scoped_acquire gil;
               (*shared)(node_id);
             });
           })
      .def("recovering_node", [](Master& m2) { return m2.GetMasterThread()->GetRecoveringNodeId(); })
      .def("all_quit", [](Master& m2) { return m2.GetMasterThread()->AllQuit(); })
      .def("heartbeats", [](Master& m2) {
        std::map<uint32_t, long long> out;
        for (const auto& kv : m2.GetMasterThread()->GetHeartBeats()) out[kv.first] = static_cast<long long>(kv.second);
        return out;
      });

  py::class_<ControlAgent>(m, "ControlAgent")
      .def(py::init<const Node&, const std::vector<Node>&, const Node&>())
      .def("start", &ControlAgent::Start, py::call_guard<py::gil_scoped_release>())
      .def("stop", &ControlAgent::Stop, py::arg("quit_heartbeat") = true, py::call_guard<py::gil_scoped_release>())
      .def("need_rollback", &ControlAgent::NeedRollBack)
      .def("clear_rollback", &ControlAgent::ClearRollBack)
      .def("rollback_count", &ControlAgent::RollBackCount)
      .def("scale", &ControlAgent::Scale);
}
