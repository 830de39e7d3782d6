// checkpoint_io.hpp — file names and the small text formats shared by the server models, the
// SVM dumper and the rollback path (reference lib/svm_dumper.hpp:47-93 DumpConfigData /
// LoadConfigData, server/util/progress_tracker.hpp:64-85 rounding).
#pragma once

#include <map>
#include <string>

namespace minips {
namespace checkpoint {

// <checkpoint_file_prefix><kind>_<my_id>   (kind: server_params, server_progress, worker_config, worker)
std::string Path(const std::string& kind);
// Checkpoints are taken every `checkpoint_interval` iterations (default 100); clocks and
// iterations are stored rounded to the nearest multiple so all parties restart aligned.
int RoundToInterval(int v);
// "<worker_id>:<iteration> ..." — the per-worker iteration map
void DumpConfigData(const std::map<int, int>& iter_map, const std::string& path = "");
std::map<int, int> LoadConfigData(const std::string& path = "");

}  // namespace checkpoint
}  // namespace minips
