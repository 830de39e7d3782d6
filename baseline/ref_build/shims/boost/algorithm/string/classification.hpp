#ifndef MINIPS_SHIM_BOOST_CLASSIFICATION_HPP_
#define MINIPS_SHIM_BOOST_CLASSIFICATION_HPP_
#include <string>
namespace boost {
struct is_any_of_pred {
  std::string set;
  bool operator()(char c) const { return set.find(c) != std::string::npos; }
};
inline is_any_of_pred is_any_of(const std::string& s) { return is_any_of_pred{s}; }
enum token_compress_mode_type { token_compress_on, token_compress_off };
}  // namespace boost
#endif
