// Test-infrastructure shim (no Boost in this image).
#pragma once
namespace boost {
template <typename T> class scoped_array {
public:
    explicit scoped_array(T *p = nullptr) : m_p(p) {}
    ~scoped_array() { delete[] m_p; }
    void reset(T *p = nullptr) { if (p != m_p) { delete[] m_p; m_p = p; } }
    T *get() const { return m_p; }
    T &operator[](long i) const { return m_p[i]; }
private:
    scoped_array(const scoped_array &); scoped_array &operator=(const scoped_array &);
    T *m_p;
};
}
