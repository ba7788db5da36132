// TEST INFRASTRUCTURE ONLY.  Included first by every translation unit of the headless
// reference build: pulls in the standard headers with their real access specifiers, then
// opens up the reference classes' private/protected members (mesh, materials, flycamera,
// intersectBox, interpolateNormal ...) so that the harness can drive and dump them.
// Access specifiers do not change layout or mangling under the Itanium ABI.
#pragma once
#include <algorithm>
#include <array>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <iomanip>
#include <iostream>
#include <limits>
#include <list>
#include <map>
#include <memory>
#include <numeric>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <tuple>
#include <unordered_map>
#include <utility>
#include <vector>
#include <complex>
#include <chrono>
#include <mutex>
#include <atomic>
#include <random>
#include <typeinfo>
#include <type_traits>
#include <iterator>
#include <new>
#define private public
#define protected public
