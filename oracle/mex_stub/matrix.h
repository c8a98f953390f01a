/* TEST INFRASTRUCTURE ONLY: MATLAB matrix.h stand-in, see mex.h in this directory. */
#include "mex.h"
