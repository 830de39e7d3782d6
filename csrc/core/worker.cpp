#include "core/worker.hpp"

namespace minips {

void WorkerThread::RegisterRecvHandle(uint32_t app_thread_id, uint32_t model_id,
                                      const std::function<void(Message&)>& recv_handle) {
  std::lock_guard<std::mutex> lk(mu_);
  recv_handles_[{app_thread_id, model_id}] = recv_handle;
}

void WorkerThread::RegisterRecvFinishHandle(uint32_t app_thread_id, uint32_t model_id,
                                            const std::function<void()>& recv_finish_handle) {
  std::lock_guard<std::mutex> lk(mu_);
  finish_handles_[{app_thread_id, model_id}] = recv_finish_handle;
}

void WorkerThread::NewRequest(uint32_t app_thread_id, uint32_t model_id, uint32_t expected_responses) {
  std::lock_guard<std::mutex> lk(mu_);
  tracker_[{app_thread_id, model_id}] = Tracker{expected_responses, 0};
}

void WorkerThread::WaitRequest(uint32_t app_thread_id, uint32_t model_id) {
  std::unique_lock<std::mutex> lk(mu_);
  const Slot s{app_thread_id, model_id};
  cond_.wait(lk, [&] { return tracker_[s].expected <= tracker_[s].current; });
}

void WorkerThread::AddResponse(uint32_t app_thread_id, uint32_t model_id, Message& msg) {
  const Slot s{app_thread_id, model_id};
  std::function<void(Message&)> recv;
  std::function<void()> finish;
  bool last = false;
  {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = recv_handles_.find(s);
    MINIPS_CHECK(it != recv_handles_.end()) << "no recv handle for thread " << app_thread_id << " model " << model_id;
    recv = it->second;
    auto& t = tracker_[s];
    last = (t.current + 1 == t.expected);
    if (last) {
      auto f = finish_handles_.find(s);
      if (f != finish_handles_.end()) finish = f->second;
    }
  }
  recv(msg);
  if (last && finish) finish();
  {
    std::lock_guard<std::mutex> lk(mu_);
    ++tracker_[s].current;
  }
  cond_.notify_all();
}

void WorkerThread::NewCheckPoint(uint32_t app_thread_id, uint32_t model_id, uint32_t expected_responses) {
  std::lock_guard<std::mutex> lk(mu_);
  ckpt_tracker_[{app_thread_id, model_id}] = Tracker{expected_responses, 0};
}

void WorkerThread::WaitCheckPoint(uint32_t app_thread_id, uint32_t model_id) {
  std::unique_lock<std::mutex> lk(mu_);
  const Slot s{app_thread_id, model_id};
  cond_.wait(lk, [&] { return ckpt_tracker_[s].expected <= ckpt_tracker_[s].current; });
}

void WorkerThread::CheckPointResponse(uint32_t app_thread_id, uint32_t model_id) {
  {
    std::lock_guard<std::mutex> lk(mu_);
    ++ckpt_tracker_[{app_thread_id, model_id}].current;
  }
  cond_.notify_all();
}

void WorkerThread::RollBackWorker() {
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (auto& kv : tracker_) kv.second.current = kv.second.expected;
    for (auto& kv : ckpt_tracker_) kv.second.current = kv.second.expected;
    finish_handles_.clear();
  }
  cond_.notify_all();
}

void WorkerThread::Update(uint32_t lost_replies_per_request) {
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (auto& kv : tracker_) {
      auto& t = kv.second;
      t.expected = t.expected > lost_replies_per_request ? t.expected - lost_replies_per_request : 0;
    }
  }
  cond_.notify_all();
}

void WorkerThread::Main() {
  while (true) {
    Message msg;
    work_queue_.WaitAndPop(&msg);
    if (msg.meta.flag == Flag::kExit) break;
    try {
      if (msg.meta.flag == Flag::kCheckpoint)
        CheckPointResponse(static_cast<uint32_t>(msg.meta.recver), static_cast<uint32_t>(msg.meta.model_id));
      else
        AddResponse(static_cast<uint32_t>(msg.meta.recver), static_cast<uint32_t>(msg.meta.model_id), msg);
    } catch (const CheckError& e) {
      MINIPS_LOG(0) << "WorkerThread " << id_ << " dropped a message: " << e.what();
    }
  }
}

}  // namespace minips
