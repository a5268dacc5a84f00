// Test-infrastructure shim (no Boost in this image): the two names spectrum.cpp uses from boost::filesystem.
#pragma once
#include <fstream>
#include <string>
namespace boost { namespace filesystem {
class path {
public:
    path() {}
    path(const char *s) : m_s(s) {}
    path(const std::string &s) : m_s(s) {}
    const std::string &string() const { return m_s; }
    const char *c_str() const { return m_s.c_str(); }
    std::string file_string() const { return m_s; }
    path filename() const { size_t k = m_s.find_last_of('/'); return path(k == std::string::npos ? m_s : m_s.substr(k + 1)); }
private:
    std::string m_s;
};
class ifstream : public std::ifstream {
public:
    ifstream() {}
    explicit ifstream(const path &p) : std::ifstream(p.c_str()) {}
};
} }
