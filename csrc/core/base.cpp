#include "core/base.hpp"

#include <cstdlib>
#include <ctime>

namespace minips {

namespace {
std::atomic<int> g_log_level{-1};
}

int LogLevel() {
  int l = g_log_level.load(std::memory_order_relaxed);
  if (l < 0) {
    const char* e = std::getenv("MINIPS_VLOG");
    l = e ? std::atoi(e) : 0;
    g_log_level.store(l);
  }
  return l;
}
void SetLogLevel(int lvl) { g_log_level.store(lvl); }

const char* FlagName(Flag f) {
  static const char* names[] = {"kExit",      "kBarrier",       "kResetWorkerInModel", "kClock",    "kAdd",
                                "kGet",       "kForceQuit",     "kCheckpoint",         "kHeartBeat", "kQuitHeartBeat",
                                "kRollBack",  "kScale",         "kScaleRollback"};
  const int i = static_cast<int>(f);
  return (i >= 0 && i < 13) ? names[i] : "kUnknown";
}

std::string Meta::DebugString() const {
  std::ostringstream os;
  os << "Meta: { sender: " << sender << ", recver: " << recver << ", model_id: " << model_id
     << ", failed_node_id: " << failed_node_id << ", flag: " << FlagName(flag) << "}";
  return os.str();
}

std::string Message::DebugString() const {
  std::ostringstream os;
  os << meta.DebugString();
  if (!data.empty()) {
    os << " Body: data_size=";
    for (const auto& d : data) os << d.size() << " ";
  }
  return os.str();
}

std::string CheckFaultTolerance(int phase) {
  static const char* what[] = {"", "", "failure detected", "failed process restarted",
                               "failed process recovered", "all nodes recovered"};
  std::ostringstream os;
  os << "[Fault Tolerance][Phase" << phase << "][" << static_cast<long long>(std::time(nullptr)) << "]";
  if (phase >= 2 && phase <= 5) os << " " << what[phase];
  MINIPS_LOG(0) << os.str();
  return os.str();
}

}  // namespace minips
