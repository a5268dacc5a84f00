// Test-infrastructure shim (no Boost in this image): the names spectrum.cpp, fstream.cpp, mmap.cpp and bitmap.cpp use
// from boost::filesystem.
#pragma once
#include <fstream>
#include <string>
#include <cstdio>
#include <cstdint>
#include <sys/stat.h>
#include <unistd.h>
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
    path extension() const { size_t k = m_s.find_last_of('.'); size_t sl = m_s.find_last_of('/');
        return path(k == std::string::npos || (sl != std::string::npos && k < sl) ? std::string() : m_s.substr(k)); }
    path &replace_extension(const std::string &e) { size_t k = m_s.find_last_of('.'); size_t sl = m_s.find_last_of('/');
        if (k != std::string::npos && (sl == std::string::npos || k > sl)) m_s.erase(k); m_s += e; return *this; }
    bool empty() const { return m_s.empty(); }
private:
    std::string m_s;
};
inline bool exists(const path &p) { struct stat st; return ::stat(p.c_str(), &st) == 0; }
inline uintmax_t file_size(const path &p) { struct stat st; return ::stat(p.c_str(), &st) == 0 ? (uintmax_t) st.st_size : 0; }
inline bool remove(const path &p) { return ::remove(p.c_str()) == 0; }
inline void resize_file(const path &p, uintmax_t n) { if (::truncate(p.c_str(), (off_t) n) != 0) {} }
class ifstream : public std::ifstream {
public:
    ifstream() {}
    explicit ifstream(const path &p) : std::ifstream(p.c_str()) {}
};
} }
