// Test-infrastructure shim (no Boost in this image).
#pragma once
#include "filesystem/fstream.hpp"
