#pragma once
#include <memory>
namespace boost {
template <typename T> class scoped_ptr {
public:
    explicit scoped_ptr(T *p = nullptr) : m_p(p) {}
    ~scoped_ptr() { delete m_p; }
    void reset(T *p = nullptr) { if (p != m_p) { delete m_p; m_p = p; } }
    T *get() const { return m_p; }
    T *operator->() const { return m_p; }
    T &operator*() const { return *m_p; }
    operator bool() const { return m_p != nullptr; }
private:
    scoped_ptr(const scoped_ptr &); scoped_ptr &operator=(const scoped_ptr &);
    T *m_p;
};
}
