#pragma once
#include <string>
#include <algorithm>
#include <cctype>
namespace boost {
inline std::string to_lower_copy(const std::string &s) {
    std::string r(s); std::transform(r.begin(), r.end(), r.begin(), [](unsigned char c) { return (char) std::tolower(c); }); return r;
}
inline std::string to_upper_copy(const std::string &s) {
    std::string r(s); std::transform(r.begin(), r.end(), r.begin(), [](unsigned char c) { return (char) std::toupper(c); }); return r;
}
inline bool ends_with(const std::string &s, const std::string &e) {
    return s.size() >= e.size() && s.compare(s.size() - e.size(), e.size(), e) == 0;
}
inline bool starts_with(const std::string &s, const std::string &e) {
    return s.size() >= e.size() && s.compare(0, e.size(), e) == 0;
}
}
