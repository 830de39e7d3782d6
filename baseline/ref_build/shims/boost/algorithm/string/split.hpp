// boost::split shim: same semantics as boost (no token compression by default: adjacent
// separators and a trailing separator yield empty tokens).
#ifndef MINIPS_SHIM_BOOST_SPLIT_HPP_
#define MINIPS_SHIM_BOOST_SPLIT_HPP_
#include <string>

#include "boost/algorithm/string/classification.hpp"
namespace boost {
template <typename Container, typename Range, typename Pred>
Container& split(Container& out, const Range& input, Pred pred, token_compress_mode_type mode = token_compress_off) {
  out.clear();
  std::string cur;
  bool last_sep = false;
  for (auto it = input.begin(); it != input.end(); ++it) {
    const char c = *it;
    if (pred(c)) {
      if (!(mode == token_compress_on && last_sep)) {
        out.push_back(typename Container::value_type(cur.begin(), cur.end()));
        cur.clear();
      }
      last_sep = true;
    } else {
      cur.push_back(c);
      last_sep = false;
    }
  }
  out.push_back(typename Container::value_type(cur.begin(), cur.end()));
  return out;
}
}  // namespace boost
#endif
