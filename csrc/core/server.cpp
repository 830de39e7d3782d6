#include "core/server.hpp"

#include <cmath>
#include <iomanip>
#include <limits>

#include "core/checkpoint_io.hpp"
#include "core/file_io.hpp"

namespace minips {

// ------------------------------------------------------------------ checkpoint paths / iteration map
namespace checkpoint {

std::string Path(const std::string& kind) {
  auto& ctx = Context::get_instance();
  return ctx.get_string("checkpoint_file_prefix") + kind + "_" + ctx.get_string("my_id");
}

int RoundToInterval(int v) {
  int interval = Context::get_instance().get_int32("checkpoint_interval");
  if (interval <= 0) interval = 100;  // the reference hard-codes 100 (progress_tracker.hpp:64-66)
  return static_cast<int>(std::lround(static_cast<double>(v) / interval)) * interval;
}

void DumpConfigData(const std::map<int, int>& iter_map, const std::string& path) {
  io::ofstream out(path.empty() ? Path("worker_config") : path, std::ios_base::out | std::ios_base::trunc);
  MINIPS_CHECK(out.is_open()) << "cannot write " << path;
  for (const auto& kv : iter_map) out << kv.first << ":" << RoundToInterval(kv.second) << " ";
}

std::map<int, int> LoadConfigData(const std::string& path) {
  std::map<int, int> m;
  io::ifstream in(path.empty() ? Path("worker_config") : path);
  if (!in.is_open()) return m;
  std::string tok;
  while (in >> tok) {
    const size_t c = tok.find(':');
    if (c == std::string::npos) continue;
    m[std::stoi(tok.substr(0, c))] = std::stoi(tok.substr(c + 1));
  }
  return m;
}

}  // namespace checkpoint

std::string ServerParamsPath() { return checkpoint::Path("server_params"); }

// ------------------------------------------------------------------ VectorStorage dump / restore
template <typename Val>
void VectorStorage<Val>::DumpTo(const std::string& path) const {
  io::ofstream out(path, std::ios_base::out | std::ios_base::trunc);
  MINIPS_CHECK(out.is_open()) << "cannot write " << path;
  out << std::setprecision(std::numeric_limits<Val>::max_digits10);
  for (size_t i = 0; i < storage_.size(); ++i)
    if (storage_[i] != Val()) out << i << ":" << storage_[i] << " ";
}

template <typename Val>
void VectorStorage<Val>::RestoreFrom(const std::string& path) {
  io::ifstream in(path);
  if (!in.is_open()) {
    MINIPS_LOG(0) << "VectorStorage::Restore: no dump at " << path;
    return;
  }
  std::fill(storage_.begin(), storage_.end(), Val());
  std::string tok;
  while (in >> tok) {
    const size_t c = tok.find(':');
    if (c == std::string::npos) continue;
    const size_t idx = std::stoull(tok.substr(0, c));
    MINIPS_CHECK_LT(idx, storage_.size());
    storage_[idx] = static_cast<Val>(std::stod(tok.substr(c + 1)));
  }
}

template class VectorStorage<int>;
template class VectorStorage<float>;
template class VectorStorage<double>;

// ------------------------------------------------------------------ ProgressTracker
void ProgressTracker::Init(const std::vector<uint32_t>& tids) {
  progresses_.clear();
  for (auto t : tids) progresses_[static_cast<int>(t)] = 0;
  min_clock_ = 0;
}

bool ProgressTracker::IsUniqueMin(int tid) const {
  auto it = progresses_.find(tid);
  if (it == progresses_.end() || it->second != min_clock_) return false;
  int at_min = 0;
  for (const auto& kv : progresses_)
    if (kv.second == min_clock_ && ++at_min > 1) return false;
  return at_min == 1;
}

int ProgressTracker::AdvanceAndGetChangedMinClock(int tid) {
  MINIPS_CHECK(CheckThreadValid(tid)) << "tid " << tid << " is not tracked";
  const bool moves_min = IsUniqueMin(tid);
  ++progresses_[tid];
  if (!moves_min) return -1;
  ++min_clock_;
  return min_clock_;
}

int ProgressTracker::GetProgress(int tid) const {
  auto it = progresses_.find(tid);
  MINIPS_CHECK(it != progresses_.end()) << "tid " << tid << " is not tracked";
  return it->second;
}

int ProgressTracker::DeleteNode(uint32_t node_id) {
  int result = -1;
  const int lo = static_cast<int>(node_id) * 1000, hi = lo + 1000;
  for (auto it = progresses_.begin(); it != progresses_.end();) {
    if (it->first >= lo && it->first < hi) {
      if (IsUniqueMin(it->first)) {
        ++min_clock_;
        result = min_clock_;
      }
      it = progresses_.erase(it);
    } else {
      ++it;
    }
  }
  return result;
}

void ProgressTracker::DumpTo(const std::string& path) const {
  io::ofstream out(path, std::ios_base::out | std::ios_base::trunc);
  MINIPS_CHECK(out.is_open()) << "cannot write " << path;
  out << "min_clock:" << checkpoint::RoundToInterval(min_clock_) << " ";
  for (const auto& kv : progresses_) out << kv.first << ":" << checkpoint::RoundToInterval(kv.second) << " ";
}

void ProgressTracker::RestoreFrom(const std::string& path) {
  io::ifstream in(path);
  if (!in.is_open()) {
    MINIPS_LOG(0) << "ProgressTracker::Restore: no dump at " << path;
    return;
  }
  std::string tok;
  while (in >> tok) {
    const size_t c = tok.find(':');
    if (c == std::string::npos) continue;
    const std::string k = tok.substr(0, c);
    const int v = std::stoi(tok.substr(c + 1));
    if (k == "min_clock") min_clock_ = v;
    else progresses_[std::stoi(k)] = v;
  }
}

void ProgressTracker::Dump() const {
  if (!Context::get_instance().get_bool("checkpoint_toggle")) return;
  DumpTo(checkpoint::Path("server_progress"));
}
void ProgressTracker::Restore() {
  if (!Context::get_instance().get_bool("checkpoint_toggle")) return;
  RestoreFrom(checkpoint::Path("server_progress"));
}

std::string ProgressTracker::DebugString() const {
  std::ostringstream os;
  os << "min_clock=" << min_clock_;
  for (const auto& kv : progresses_) os << " " << kv.first << ":" << kv.second;
  return os.str();
}

// ------------------------------------------------------------------ PendingBuffer
std::vector<Message> PendingBuffer::Pop(int clock) {
  std::vector<Message> out;
  auto it = buffer_.find(clock);
  if (it != buffer_.end()) {
    out = std::move(it->second);
    buffer_.erase(it);
  }
  return out;
}

std::vector<Message> PendingBuffer::PopAll() {
  std::vector<Message> out;
  for (auto& kv : buffer_)
    for (auto& m : kv.second) out.push_back(std::move(m));
  buffer_.clear();
  return out;
}

int PendingBuffer::Size(int progress) const {
  auto it = buffer_.find(progress);
  return it == buffer_.end() ? 0 : static_cast<int>(it->second.size());
}

// ------------------------------------------------------------------ ModelBase
void ModelBase::ResetWorker(Message& msg) {
  MINIPS_CHECK_EQ(msg.data.size(), 1u);
  SArray<uint32_t> tids(msg.data[0]);
  progress_tracker_.Init(std::vector<uint32_t>(tids.begin(), tids.end()));
  Message ack;
  ack.meta.recver = msg.meta.sender;
  ack.meta.sender = msg.meta.recver;
  ack.meta.flag = Flag::kResetWorkerInModel;
  ack.meta.model_id = model_id_;
  reply_queue_->Push(ack);
}

void ModelBase::Dump(Message& msg) {
  storage_->Dump();
  progress_tracker_.Dump();
  if (Context::get_instance().get_bool("checkpoint_toggle"))
    checkpoint::DumpConfigData(Context::get_instance().get_iteration_map());
  Message ack;
  ack.meta.recver = msg.meta.sender;
  ack.meta.sender = msg.meta.recver;
  ack.meta.flag = Flag::kCheckpoint;
  ack.meta.model_id = model_id_;
  reply_queue_->Push(ack);
}

void ModelBase::Restore() {
  storage_->Restore();
  progress_tracker_.Restore();
  AfterRestore();
}

void ModelBase::Update(int failed_node_id, const std::vector<Node>& /*nodes*/, const Range& range) {
  if (failed_node_id >= 0) progress_tracker_.DeleteNode(static_cast<uint32_t>(failed_node_id));
  storage_->Update(range);
}

// ------------------------------------------------------------------ BSP
void BSPModel::Clock(Message& msg) {
  const int new_min = progress_tracker_.AdvanceAndGetChangedMinClock(msg.meta.sender);
  if (new_min == -1) return;
  // clock edge: buffered adds become visible, then the gets that waited for them are served
  for (auto& add : add_buffer_) storage_->Add(add);
  add_buffer_.clear();
  for (auto& get : get_buffer_) reply_queue_->Push(storage_->Get(get));
  get_buffer_.clear();
  storage_->FinishIter();
}

void BSPModel::Add(Message& msg) {
  MINIPS_CHECK(progress_tracker_.CheckThreadValid(msg.meta.sender));
  const int progress = progress_tracker_.GetProgress(msg.meta.sender);
  MINIPS_CHECK(progress == progress_tracker_.GetMinClock())
      << "BSP add from a worker ahead of the barrier: progress " << progress << " min "
      << progress_tracker_.GetMinClock();
  add_buffer_.push_back(msg);
}

void BSPModel::Get(Message& msg) {
  MINIPS_CHECK(progress_tracker_.CheckThreadValid(msg.meta.sender));
  const int progress = progress_tracker_.GetProgress(msg.meta.sender);
  const int min_clock = progress_tracker_.GetMinClock();
  if (progress == min_clock) {
    reply_queue_->Push(storage_->Get(msg));
  } else {
    MINIPS_CHECK(progress == min_clock + 1) << "BSP get more than one clock ahead: " << progress << " vs " << min_clock;
    get_buffer_.push_back(msg);
  }
}

void BSPModel::AfterRestore() {
  add_buffer_.clear();
  for (auto& get : get_buffer_) reply_queue_->Push(storage_->Get(get));
  get_buffer_.clear();
}

// ------------------------------------------------------------------ SSP
SSPModel::SSPModel(uint32_t model_id, std::unique_ptr<AbstractStorage>&& storage, int staleness,
                   ThreadsafeQueue<Message>* reply_queue)
    : ModelBase(model_id, std::move(storage), reply_queue), staleness_(staleness) {
  if (Context::get_instance().get_bool("use_weight_file")) {
    storage_->Restore();
    progress_tracker_.Restore();
  }
}

void SSPModel::Clock(Message& msg) {
  const int new_min = progress_tracker_.AdvanceAndGetChangedMinClock(msg.meta.sender);
  if (new_min != -1) Flush(new_min);
}

void SSPModel::Flush(int updated_min_clock) {
  for (auto& req : buffer_.Pop(updated_min_clock)) reply_queue_->Push(storage_->Get(req));
  storage_->FinishIter();
}

void SSPModel::Add(Message& msg) { storage_->Add(msg); }

void SSPModel::Get(Message& msg) {
  MINIPS_CHECK(progress_tracker_.CheckThreadValid(msg.meta.sender));
  const int progress = progress_tracker_.GetProgress(msg.meta.sender);
  const int min_clock = progress_tracker_.GetMinClock();
  if (progress > min_clock + staleness_) buffer_.Push(progress - staleness_, msg);
  else reply_queue_->Push(storage_->Get(msg));
}

void SSPModel::ResetWorker(Message& msg) {
  if (Context::get_instance().get_bool("use_weight_file")) {
    // keep the restored progress (ssp_model.cpp:98-103), still acknowledge
    Message ack;
    ack.meta.recver = msg.meta.sender;
    ack.meta.sender = msg.meta.recver;
    ack.meta.flag = Flag::kResetWorkerInModel;
    ack.meta.model_id = model_id_;
    reply_queue_->Push(ack);
    return;
  }
  ModelBase::ResetWorker(msg);
}

void SSPModel::AfterRestore() {
  for (auto& req : buffer_.PopAll()) reply_queue_->Push(storage_->Get(req));
  storage_->FinishIter();
}

// ------------------------------------------------------------------ ASP
void ASPModel::Clock(Message& msg) { progress_tracker_.AdvanceAndGetChangedMinClock(msg.meta.sender); }
void ASPModel::Add(Message& msg) {
  MINIPS_CHECK(progress_tracker_.CheckThreadValid(msg.meta.sender));
  storage_->Add(msg);
}
void ASPModel::Get(Message& msg) {
  MINIPS_CHECK(progress_tracker_.CheckThreadValid(msg.meta.sender));
  reply_queue_->Push(storage_->Get(msg));
}

// ------------------------------------------------------------------ ServerThread
void ServerThread::RegisterModel(uint32_t model_id, std::unique_ptr<AbstractModel>&& model) {
  std::lock_guard<std::mutex> lk(mu_);
  MINIPS_CHECK(models_.find(model_id) == models_.end()) << "model " << model_id << " registered twice";
  models_[model_id] = std::move(model);
}

AbstractModel* ServerThread::GetModel(uint32_t model_id) {
  std::lock_guard<std::mutex> lk(mu_);
  auto it = models_.find(model_id);
  MINIPS_CHECK(it != models_.end()) << "unknown model " << model_id;
  return it->second.get();
}

void ServerThread::UpdateModel(int failed_node_id, const std::vector<Node>& nodes, const Range& range) {
  std::lock_guard<std::mutex> lk(mu_);
  for (auto& kv : models_) kv.second->Update(failed_node_id, nodes, range);
}

void ServerThread::RollbackModel() {
  std::lock_guard<std::mutex> lk(mu_);
  for (auto& kv : models_) kv.second->Restore();
}

void ServerThread::Main() {
  while (true) {
    Message msg;
    work_queue_.WaitAndPop(&msg);
    if (msg.meta.flag == Flag::kExit) break;
    std::lock_guard<std::mutex> lk(mu_);
    auto it = models_.find(static_cast<uint32_t>(msg.meta.model_id));
    if (it == models_.end()) {
      MINIPS_LOG(0) << "ServerThread " << id_ << ": unknown model in " << msg.DebugString();
      continue;
    }
    AbstractModel* model = it->second.get();
    try {
      switch (msg.meta.flag) {
        case Flag::kClock: model->Clock(msg); break;
        case Flag::kAdd: model->Add(msg); break;
        case Flag::kGet: model->Get(msg); break;
        case Flag::kResetWorkerInModel: model->ResetWorker(msg); break;
        case Flag::kCheckpoint: model->Dump(msg); break;
        default: MINIPS_LOG(0) << "ServerThread " << id_ << ": unexpected " << msg.DebugString();
      }
    } catch (const CheckError& e) {
      MINIPS_LOG(0) << "ServerThread " << id_ << " dropped a message: " << e.what();
    }
  }
}

}  // namespace minips
