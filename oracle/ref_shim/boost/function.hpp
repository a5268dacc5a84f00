// Test-infrastructure shim (no Boost in this image): boost::function as std::function.
#pragma once
#include <functional>
namespace boost { template <typename Sig> using function = std::function<Sig>; }
