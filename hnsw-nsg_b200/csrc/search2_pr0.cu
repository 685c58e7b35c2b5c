// search kernel v2, pool form PR=0 (see search2_impl.cuh)
#define SEARCH2_PR 0
#include "search2_impl.cuh"
