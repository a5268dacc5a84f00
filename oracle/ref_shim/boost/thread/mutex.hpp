// Test-infrastructure shim (no Boost in this image).
#pragma once
#include <mutex>
namespace boost {
class mutex : public std::mutex {
public:
    typedef std::unique_lock<std::mutex> scoped_lock;
};
}
