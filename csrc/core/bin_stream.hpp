// bin_stream.hpp — byte-stream serialisation used by the block-assigner protocol and for
// shipping SArrays (reference base/serialization.hpp:31-131, .cpp:22-136: BinStream with
// operator<< / operator>> for PODs, strings, vectors, maps, pairs, smart pointers, nested
// streams, and types exposing serialize()/deserialize()).
#pragma once

#include <map>
#include <memory>
#include <string>
#include <type_traits>
#include <unordered_map>
#include <utility>
#include <vector>

#include "core/base.hpp"

namespace minips {

class BinStream {
 public:
  BinStream() = default;
  explicit BinStream(size_t reserve) { buf_.reserve(reserve); }
  BinStream(const char* src, size_t n) : buf_(src, src + n) {}
  explicit BinStream(std::vector<char> v) : buf_(std::move(v)) {}

  size_t size() const { return buf_.size() - front_; }
  bool empty() const { return size() == 0; }
  void clear() {
    buf_.clear();
    front_ = 0;
  }
  void seek(size_t pos) { front_ = pos; }
  void purge() {  // drop the consumed prefix
    buf_.erase(buf_.begin(), buf_.begin() + front_);
    front_ = 0;
  }
  void resize(size_t n) { buf_.resize(front_ + n); }
  void append(const BinStream& o) { push_back_bytes(o.get_remained_buffer(), o.size()); }
  void push_back_bytes(const char* src, size_t n) { buf_.insert(buf_.end(), src, src + n); }
  void* pop_front_bytes(size_t n) {
    MINIPS_CHECK(front_ + n <= buf_.size()) << "BinStream underflow: need " << n << " have " << size();
    void* p = buf_.data() + front_;
    front_ += n;
    return p;
  }
  const char* get_remained_buffer() const { return buf_.data() + front_; }
  char* get_buffer() { return buf_.data(); }
  const std::vector<char>& get_buffer_vector() const { return buf_; }
  std::string to_string() const { return std::string(get_remained_buffer(), size()); }

 private:
  std::vector<char> buf_;
  size_t front_ = 0;
};

namespace detail {
template <typename T, typename = void>
struct has_serialize : std::false_type {};
template <typename T>
struct has_serialize<T, std::void_t<decltype(std::declval<const T&>().serialize(std::declval<BinStream&>()))>>
    : std::true_type {};
template <typename T, typename = void>
struct has_deserialize : std::false_type {};
template <typename T>
struct has_deserialize<T, std::void_t<decltype(std::declval<T&>().deserialize(std::declval<BinStream&>()))>>
    : std::true_type {};
}  // namespace detail

// ---- PODs and user types
template <typename T>
std::enable_if_t<std::is_trivially_copyable<T>::value && !detail::has_serialize<T>::value, BinStream&>
operator<<(BinStream& s, const T& x) {
  s.push_back_bytes(reinterpret_cast<const char*>(&x), sizeof(T));
  return s;
}
template <typename T>
std::enable_if_t<std::is_trivially_copyable<T>::value && !detail::has_deserialize<T>::value, BinStream&>
operator>>(BinStream& s, T& x) {
  std::memcpy(&x, s.pop_front_bytes(sizeof(T)), sizeof(T));
  return s;
}
template <typename T>
std::enable_if_t<detail::has_serialize<T>::value, BinStream&> operator<<(BinStream& s, const T& x) {
  x.serialize(s);
  return s;
}
template <typename T>
std::enable_if_t<detail::has_deserialize<T>::value, BinStream&> operator>>(BinStream& s, T& x) {
  x.deserialize(s);
  return s;
}

// ---- strings
inline BinStream& operator<<(BinStream& s, const std::string& x) {
  s << static_cast<uint64_t>(x.size());
  s.push_back_bytes(x.data(), x.size());
  return s;
}
inline BinStream& operator>>(BinStream& s, std::string& x) {
  uint64_t n = 0;
  s >> n;
  const char* p = static_cast<const char*>(s.pop_front_bytes(n));
  x.assign(p, n);
  return s;
}
inline BinStream& operator<<(BinStream& s, const char* x) { return s << std::string(x); }

// ---- forward declarations so containers of containers resolve
template <typename A, typename B>
BinStream& operator<<(BinStream& s, const std::pair<A, B>& p);
template <typename A, typename B>
BinStream& operator>>(BinStream& s, std::pair<A, B>& p);
template <typename T>
BinStream& operator<<(BinStream& s, const std::vector<T>& v);
template <typename T>
BinStream& operator>>(BinStream& s, std::vector<T>& v);

template <typename A, typename B>
BinStream& operator<<(BinStream& s, const std::pair<A, B>& p) {
  return s << p.first << p.second;
}
template <typename A, typename B>
BinStream& operator>>(BinStream& s, std::pair<A, B>& p) {
  return s >> p.first >> p.second;
}

template <typename T>
BinStream& operator<<(BinStream& s, const std::vector<T>& v) {
  s << static_cast<uint64_t>(v.size());
  for (const auto& e : v) s << e;
  return s;
}
template <typename T>
BinStream& operator>>(BinStream& s, std::vector<T>& v) {
  uint64_t n = 0;
  s >> n;
  v.clear();
  v.reserve(n);
  for (uint64_t i = 0; i < n; ++i) {
    T e;
    s >> e;
    v.push_back(std::move(e));
  }
  return s;
}
inline BinStream& operator<<(BinStream& s, const std::vector<bool>& v) {
  s << static_cast<uint64_t>(v.size());
  for (bool b : v) s << static_cast<char>(b ? 1 : 0);
  return s;
}
inline BinStream& operator>>(BinStream& s, std::vector<bool>& v) {
  uint64_t n = 0;
  s >> n;
  v.assign(n, false);
  for (uint64_t i = 0; i < n; ++i) {
    char c = 0;
    s >> c;
    v[i] = c != 0;
  }
  return s;
}

template <typename K, typename V>
BinStream& operator<<(BinStream& s, const std::map<K, V>& m) {
  s << static_cast<uint64_t>(m.size());
  for (const auto& kv : m) s << kv.first << kv.second;
  return s;
}
template <typename K, typename V>
BinStream& operator>>(BinStream& s, std::map<K, V>& m) {
  uint64_t n = 0;
  s >> n;
  m.clear();
  for (uint64_t i = 0; i < n; ++i) {
    K k;
    V v;
    s >> k >> v;
    m.emplace(std::move(k), std::move(v));
  }
  return s;
}
template <typename K, typename V>
BinStream& operator<<(BinStream& s, const std::unordered_map<K, V>& m) {
  s << static_cast<uint64_t>(m.size());
  for (const auto& kv : m) s << kv.first << kv.second;
  return s;
}
template <typename K, typename V>
BinStream& operator>>(BinStream& s, std::unordered_map<K, V>& m) {
  uint64_t n = 0;
  s >> n;
  m.clear();
  for (uint64_t i = 0; i < n; ++i) {
    K k;
    V v;
    s >> k >> v;
    m.emplace(std::move(k), std::move(v));
  }
  return s;
}

template <typename T>
BinStream& operator<<(BinStream& s, const std::shared_ptr<T>& p) {
  return s << *p;
}
template <typename T>
BinStream& operator>>(BinStream& s, std::shared_ptr<T>& p) {
  p = std::make_shared<T>();
  return s >> *p;
}
template <typename T>
BinStream& operator<<(BinStream& s, const std::unique_ptr<T>& p) {
  return s << *p;
}
template <typename T>
BinStream& operator>>(BinStream& s, std::unique_ptr<T>& p) {
  p.reset(new T);
  return s >> *p;
}

// nested stream: length-prefixed
inline BinStream& operator<<(BinStream& s, const BinStream& inner) {
  s << static_cast<uint64_t>(inner.size());
  s.push_back_bytes(inner.get_remained_buffer(), inner.size());
  return s;
}
inline BinStream& operator>>(BinStream& s, BinStream& inner) {
  uint64_t n = 0;
  s >> n;
  inner.clear();
  inner.push_back_bytes(static_cast<const char*>(s.pop_front_bytes(n)), n);
  return s;
}

// SArray (sarray.h:270-282)
template <typename V>
BinStream& operator<<(BinStream& s, const SArray<V>& a) {
  s << static_cast<uint64_t>(a.size());
  s.push_back_bytes(reinterpret_cast<const char*>(a.data()), a.size() * sizeof(V));
  return s;
}
template <typename V>
BinStream& operator>>(BinStream& s, SArray<V>& a) {
  uint64_t n = 0;
  s >> n;
  a.CopyFrom(static_cast<const V*>(s.pop_front_bytes(n * sizeof(V))), n);
  return s;
}

}  // namespace minips
