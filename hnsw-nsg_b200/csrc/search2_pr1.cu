// search kernel v2, pool form PR=1 (see search2_impl.cuh)
#define SEARCH2_PR 1
#include "search2_impl.cuh"
