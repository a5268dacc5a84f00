// Test-infrastructure shim (no Boost in this image): boost::bind / _1.. as std::bind and its placeholders.
#pragma once
#include <functional>
namespace boost { using std::bind; using std::ref; using std::cref; }
using std::placeholders::_1; using std::placeholders::_2; using std::placeholders::_3; using std::placeholders::_4;
