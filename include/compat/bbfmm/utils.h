// aniso_b200 drop-in header: replaces bbfmm/utils.h of lowrank/aniso (constants and typedefs only; reference
// bbfmm/utils.h:46-81).  Callers written against the reference (main.cpp, AnisoWrapper.cpp) compile unchanged with
// -I<repo>/include/compat and link against libaniso_b200.so.
#ifndef ANISO_B200_COMPAT_UTILS_H
#define ANISO_B200_COMPAT_UTILS_H

#include <algorithm>
#include <cassert>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <iomanip>
#include <iostream>
#include <queue>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#ifdef RUN_OMP
#include "omp.h"
#endif

#define EPS 1e-12
#define DIM 2
#define SQR(X) ((X) * (X))
#define RUN(s, func) \
    { func; }
#define CHUNKSIZE 80

typedef int index_t;
typedef double scalar_t;
typedef bool bool_t;

using std::max;
using std::min;
using std::queue;
using std::sqrt;
using std::unordered_map;
using std::unordered_set;
using std::vector;

#endif
