#ifndef MINIPS_SHIM_BOOST_PREDICATE_HPP_
#define MINIPS_SHIM_BOOST_PREDICATE_HPP_
#include <string>
namespace boost {
inline bool starts_with(const std::string& s, const std::string& prefix) {
  return s.size() >= prefix.size() && s.compare(0, prefix.size(), prefix) == 0;
}
inline bool ends_with(const std::string& s, const std::string& suffix) {
  return s.size() >= suffix.size() && s.compare(s.size() - suffix.size(), suffix.size(), suffix) == 0;
}
}  // namespace boost
#endif
