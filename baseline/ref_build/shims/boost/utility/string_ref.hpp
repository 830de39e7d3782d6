// boost/utility/string_ref.hpp shim — non-owning string view with the members the reference uses.
#ifndef MINIPS_SHIM_STRING_REF_HPP_
#define MINIPS_SHIM_STRING_REF_HPP_
#include <algorithm>
#include <cstring>
#include <ostream>
#include <string>
namespace boost {
class string_ref {
 public:
  typedef const char* const_iterator;
  typedef const_iterator iterator;
  static const size_t npos = static_cast<size_t>(-1);
  string_ref() : p_(nullptr), n_(0) {}
  string_ref(const char* s) : p_(s), n_(s ? std::strlen(s) : 0) {}
  string_ref(const char* s, size_t n) : p_(s), n_(n) {}
  string_ref(const std::string& s) : p_(s.data()), n_(s.size()) {}
  const char* data() const { return p_; }
  size_t size() const { return n_; }
  size_t length() const { return n_; }
  bool empty() const { return n_ == 0; }
  const_iterator begin() const { return p_; }
  const_iterator end() const { return p_ + n_; }
  const char& operator[](size_t i) const { return p_[i]; }
  const char& front() const { return p_[0]; }
  const char& back() const { return p_[n_ - 1]; }
  void clear() { n_ = 0; }
  void remove_prefix(size_t n) { p_ += n; n_ -= n; }
  void remove_suffix(size_t n) { n_ -= n; }
  string_ref substr(size_t pos, size_t n = npos) const {
    if (pos > n_) pos = n_;
    return string_ref(p_ + pos, std::min(n, n_ - pos));
  }
  size_t find(char c, size_t pos = 0) const {
    for (size_t i = pos; i < n_; ++i)
      if (p_[i] == c) return i;
    return npos;
  }
  size_t find(string_ref s) const {
    if (s.n_ > n_) return npos;
    for (size_t i = 0; i + s.n_ <= n_; ++i)
      if (std::memcmp(p_ + i, s.p_, s.n_) == 0) return i;
    return npos;
  }
  std::string to_string() const { return std::string(p_, n_); }
  explicit operator std::string() const { return to_string(); }
  int compare(string_ref o) const {
    const int c = std::memcmp(p_, o.p_, std::min(n_, o.n_));
    return c ? c : (n_ == o.n_ ? 0 : (n_ < o.n_ ? -1 : 1));
  }

 private:
  const char* p_;
  size_t n_;
};
inline bool operator==(string_ref a, string_ref b) { return a.size() == b.size() && a.compare(b) == 0; }
inline bool operator!=(string_ref a, string_ref b) { return !(a == b); }
inline std::ostream& operator<<(std::ostream& os, string_ref s) {
  return os.write(s.data(), static_cast<std::streamsize>(s.size()));
}
}  // namespace boost
#endif
