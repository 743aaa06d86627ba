#pragma once
#include <string>
#include <vector>
namespace boost {
  struct is_any_of {
    std::string chars;
    explicit is_any_of(const char *c) : chars(c) {}
    bool operator()(char c) const { return chars.find(c) != std::string::npos; }
  };
  template <typename Seq, typename Pred>
  void split(Seq &out, const std::string &in, Pred pred) {
    out.clear();
    std::string cur;
    for (char c : in) {
      if (pred(c)) { out.push_back(cur); cur.clear(); } else cur.push_back(c);
    }
    out.push_back(cur);
  }
}  // namespace boost
