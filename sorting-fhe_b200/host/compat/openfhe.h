// openfhe.h -- forwarding header: lets the reference's UNMODIFIED sources (src/sort_algo.h, comparison.*, sign.*,
// rotation.h, encryption.*, k-way/*, mehp24/*) compile against the B200 engine.  Put this directory before any real
// OpenFHE installation on the include path (-I sorting-fhe_b200/host/compat) and link libckks_b200.so.
// OpenFHE's umbrella header also leaks a set of standard headers and the `usint` typedef the reference relies on.
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <complex>
#include <cstdint>
#include <fstream>
#include <functional>
#include <iomanip>
#include <iostream>
#include <map>
#include <memory>
#include <set>
#include <sstream>
#include <string>
#include <vector>
typedef uint32_t usint;
#include "../openfhe_shim.h"
