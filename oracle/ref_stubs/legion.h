// Stand-in for Legion's header when reference HEADERS (flexflow/batch_config.h and what it includes) are compiled into
// oracle/_ref/: the batch descriptors only name Legion::Future in type aliases.  Same stub as tests/test_abi.py uses.
#pragma once
#include <cstddef>
#include <iostream>
#include <string>
namespace Legion {
struct Future {};
typedef long long coord_t;
struct Memory {};
} // namespace Legion
