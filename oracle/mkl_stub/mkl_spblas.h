/* mkl_spblas.h -- stand-in, see mkl.h in this directory.  TEST INFRASTRUCTURE ONLY. */
#include "mkl.h"
