// partition.hpp — how keys map to server threads.
//
// AbstractIdMapper          base/abstract_id_mapper.hpp:7-12
// AbstractPartitionManager  base/abstract_partition_manager.hpp:14-49
// RangePartitionManager     base/range_partition_manager.hpp:14-76 (sorted-key slicing into
//                           zero-copy segments, empty servers skipped, keys outside every
//                           range dropped, k values per key via the vals/keys size ratio)
#pragma once

#include <algorithm>
#include <utility>
#include <vector>

#include "core/base.hpp"

namespace minips {

class AbstractIdMapper {
 public:
  virtual ~AbstractIdMapper() = default;
  virtual uint32_t GetNodeIdForThread(uint32_t tid) = 0;
};

class AbstractPartitionManager {
 public:
  using Keys = SArray<Key>;
  using KVPairs = std::pair<SArray<Key>, SArray<double>>;

  explicit AbstractPartitionManager(const std::vector<uint32_t>& server_thread_ids, int master_node_id = -1)
      : server_thread_ids_(server_thread_ids), master_node_id_(master_node_id) {}
  virtual ~AbstractPartitionManager() = default;

  size_t GetNumServers() const { return server_thread_ids_.size(); }
  const std::vector<uint32_t>& GetServerThreadIds() const { return server_thread_ids_; }
  int GetMasterNodeId() const { return master_node_id_; }

  virtual void Slice(const Keys& keys, std::vector<std::pair<int, Keys>>* sliced) const = 0;
  virtual void Slice(const KVPairs& kvs, std::vector<std::pair<int, KVPairs>>* sliced) const = 0;
  virtual void Update(const std::vector<Range>& ranges, const std::vector<uint32_t>& server_thread_ids) = 0;

 protected:
  std::vector<uint32_t> server_thread_ids_;
  int master_node_id_;
};

class RangePartitionManager : public AbstractPartitionManager {
 public:
  RangePartitionManager(const std::vector<uint32_t>& server_thread_ids, const std::vector<Range>& ranges,
                        int master_node_id = -1)
      : AbstractPartitionManager(server_thread_ids, master_node_id), ranges_(ranges) {
    MINIPS_CHECK_EQ(ranges_.size(), server_thread_ids_.size());
  }

  const std::vector<Range>& GetRanges() const { return ranges_; }

  // One pass over the sorted keys: for server i the segment is [pos, lower_bound(end_i)).
  void Slice(const Keys& keys, std::vector<std::pair<int, Keys>>* sliced) const override {
    sliced->clear();
    if (keys.empty() || ranges_.empty()) return;
    const Key* first = keys.begin();
    const Key* last = keys.end();
    const Key* pos = std::lower_bound(first, last, static_cast<Key>(ranges_.front().begin()));
    for (size_t i = 0; i < ranges_.size() && pos != last; ++i) {
      const Key* stop = ranges_[i].end() > 0xFFFFFFFFull
                            ? last
                            : std::lower_bound(pos, last, static_cast<Key>(ranges_[i].end()));
      if (stop != pos)
        sliced->emplace_back(static_cast<int>(server_thread_ids_[i]), keys.segment(pos - first, stop - first));
      pos = stop;
    }
  }

  void Slice(const KVPairs& kvs, std::vector<std::pair<int, KVPairs>>* sliced) const override {
    sliced->clear();
    const Keys& keys = kvs.first;
    if (keys.empty() || ranges_.empty()) return;
    MINIPS_CHECK(kvs.second.size() % keys.size() == 0) << "values must be a multiple of keys";
    const size_t ratio = kvs.second.size() / keys.size();
    const Key* first = keys.begin();
    const Key* last = keys.end();
    const Key* pos = std::lower_bound(first, last, static_cast<Key>(ranges_.front().begin()));
    for (size_t i = 0; i < ranges_.size() && pos != last; ++i) {
      const Key* stop = ranges_[i].end() > 0xFFFFFFFFull
                            ? last
                            : std::lower_bound(pos, last, static_cast<Key>(ranges_[i].end()));
      if (stop != pos) {
        const size_t b = pos - first, e = stop - first;
        sliced->emplace_back(static_cast<int>(server_thread_ids_[i]),
                             KVPairs(keys.segment(b, e), kvs.second.segment(b * ratio, e * ratio)));
      }
      pos = stop;
    }
  }

  void Update(const std::vector<Range>& ranges, const std::vector<uint32_t>& server_thread_ids) override {
    MINIPS_CHECK_EQ(ranges.size(), server_thread_ids.size());
    ranges_ = ranges;
    server_thread_ids_ = server_thread_ids;
  }

 private:
  std::vector<Range> ranges_;
};

}  // namespace minips
